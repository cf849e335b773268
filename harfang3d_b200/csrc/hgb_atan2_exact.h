// atan2 as the reference computes it — bit for bit.
//
// ToEuler (foundation/matrix3.cpp:91-137) calls the C library's DOUBLE atan2 on float arguments and narrows the result to float. The
// host libm (glibc 2.39, IBM Accurate Mathematical Library) returns the correctly rounded double; CUDA's double atan2 is within 2 ulp.
// After the narrowing both give the same float unless a float rounding boundary (the midpoint of two adjacent floats) lies between
// the two doubles, i.e. within ~3 double-ulps of CUDA's result — about 1.5e-8 of all evaluations. hgb_atan2f_ref() therefore takes
// the library result whenever it is more than 8 double-ulps away from every float midpoint, and otherwise recomputes the angle in
// double-double arithmetic (~100 bits: table of atan(k/16) + Taylor series on the reduced argument), rounds it ONCE to double — the
// value a correctly rounded libm returns — and narrows that. Both roundings of the reference are reproduced, including the tie case
// in which the double result IS the float midpoint.
//
// Plain C++ with explicit fma(): compiles for the device (nvcc, -fmad=false) and for the host (g++ -ffp-contract=off), where
// oracle/exhaustive_atan2.c checks it against the libm the reference links on ~10^9 sampled and near-midpoint inputs.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define HGB_HD __host__ __device__
#else
#define HGB_HD
#endif

namespace hgb_exact {

struct dd {
	double hi, lo;
};

HGB_HD inline dd two_sum(double a, double b) {
	const double s = a + b, bb = s - a;
	return dd{s, (a - (s - bb)) + (b - bb)};
}
HGB_HD inline dd quick_two_sum(double a, double b) { // |a| >= |b|
	const double s = a + b;
	return dd{s, b - (s - a)};
}
HGB_HD inline dd two_prod(double a, double b) {
	const double p = a * b;
	return dd{p, fma(a, b, -p)};
}
HGB_HD inline dd dd_add(dd a, dd b) {
	dd s = two_sum(a.hi, b.hi);
	const dd t = two_sum(a.lo, b.lo);
	s.lo += t.hi;
	s = quick_two_sum(s.hi, s.lo);
	s.lo += t.lo;
	return quick_two_sum(s.hi, s.lo);
}
HGB_HD inline dd dd_neg(dd a) { return dd{-a.hi, -a.lo}; }
HGB_HD inline dd dd_mul(dd a, dd b) {
	dd p = two_prod(a.hi, b.hi);
	p.lo += a.hi * b.lo + a.lo * b.hi;
	return quick_two_sum(p.hi, p.lo);
}
HGB_HD inline dd dd_div(dd a, dd b) { // three-step long division
	const double q1 = a.hi / b.hi;
	dd r = dd_add(a, dd_neg(dd_mul(b, dd{q1, 0.0})));
	const double q2 = r.hi / b.hi;
	r = dd_add(r, dd_neg(dd_mul(b, dd{q2, 0.0})));
	const double q3 = r.hi / b.hi;
	const dd q = quick_two_sum(q1, q2);
	return dd_add(q, dd{q3, 0.0});
}

// atan2(y, x) for finite non-zero doubles that hold float values, as a normalised double-double (relative error ~2^-98)
HGB_HD inline dd atan2_dd(double y, double x) {
	const double a_hi[17] = {0x0.0p+0, 0x1.ff55bb72cfdeap-5, 0x1.fd5ba9aac2f6ep-4, 0x1.7b97b4bce5b02p-3, 0x1.f5b75f92c80ddp-3, 0x1.362773707ebccp-2,
		0x1.6f61941e4def1p-2, 0x1.a64eec3cc23fdp-2, 0x1.dac670561bb4fp-2, 0x1.0657e94db30d0p-1, 0x1.1e00babdefeb4p-1, 0x1.345f01cce37bbp-1,
		0x1.4978fa3269ee1p-1, 0x1.5d58987169b18p-1, 0x1.700a7c5784634p-1, 0x1.819d0b7158a4dp-1, 0x1.921fb54442d18p-1};
	const double a_lo[17] = {0x0.0p+0, -0x1.c934d86d23f1dp-60, -0x1.cd37686760c17p-59, 0x1.347b0b4f881cap-58, 0x1.8ab6e3cf7afbdp-57, -0x1.963a544b672d8p-57,
		-0x1.c63aae6f6e918p-56, -0x1.24dec1b50b7ffp-56, 0x1.a2b7f222f65e2p-56, -0x1.d5b495f6349e6p-56, -0x1.928df287a668fp-58, 0x1.1021137c71102p-55,
		0x1.2419a87f2a458p-56, 0x1.0028e4bc5e7cap-57, -0x1.8c34d25aadef6p-56, -0x1.bf76229d3b917p-56, 0x1.1a62633145c07p-55};
	// 1/3, 1/5, ... 1/23
	const double c_hi[11] = {0x1.5555555555555p-2, 0x1.999999999999ap-3, 0x1.2492492492492p-3, 0x1.c71c71c71c71cp-4, 0x1.745d1745d1746p-4, 0x1.3b13b13b13b14p-4,
		0x1.1111111111111p-4, 0x1.e1e1e1e1e1e1ep-5, 0x1.af286bca1af28p-5, 0x1.8618618618618p-5, 0x1.642c8590b2164p-5};
	const double c_lo[11] = {0x1.5555555555555p-56, -0x1.999999999999ap-57, 0x1.2492492492492p-57, 0x1.c71c71c71c71cp-58, -0x1.745d1745d1746p-59,
		-0x1.3b13b13b13b14p-58, 0x1.1111111111111p-60, 0x1.e1e1e1e1e1e1ep-61, 0x1.af286bca1af28p-59, 0x1.8618618618618p-59, 0x1.642c8590b2164p-60};
	const dd pi = {0x1.921fb54442d18p+1, 0x1.1a62633145c07p-53}, pio2 = {0x1.921fb54442d18p+0, 0x1.1a62633145c07p-54};

	const double ax = fabs(x), ay = fabs(y);
	const bool swap = ay > ax;
	const dd q = swap ? dd_div(dd{ax, 0.0}, dd{ay, 0.0}) : dd_div(dd{ay, 0.0}, dd{ax, 0.0}); // in [0, 1]
	const int k = int(q.hi * 16.0 + 0.5);
	dd t = q;
	if (k > 0) { // atan(q) = atan(c) + atan((q - c) / (1 + q c)), c = k/16 (exact)
		const dd c = {double(k) * 0.0625, 0.0};
		t = dd_div(dd_add(q, dd_neg(c)), dd_add(dd{1.0, 0.0}, dd_mul(q, c)));
	}
	// atan(t) = t (1 - s/3 + s^2/5 - ...), s = t^2, |t| <= 1/31: eleven terms reach 2^-110
	const dd s = dd_mul(t, t);
	dd p = {c_hi[10], c_lo[10]};
	for (int n = 9; n >= 0; --n)
		p = dd_add(dd{c_hi[n], c_lo[n]}, dd_neg(dd_mul(p, s)));
	p = dd_add(dd{1.0, 0.0}, dd_neg(dd_mul(p, s)));
	dd r = dd_add(dd{a_hi[k], a_lo[k]}, dd_mul(t, p));
	if (swap)
		r = dd_add(pio2, dd_neg(r));
	if (x < 0.0)
		r = dd_add(pi, dd_neg(r));
	return y < 0.0 ? dd_neg(r) : r;
}

HGB_HD inline uint64_t d2u(double d) {
#if defined(__CUDA_ARCH__)
	return uint64_t(__double_as_longlong(d));
#else
	uint64_t u;
	memcpy(&u, &d, 8);
	return u;
#endif
}

// true when narrowing `d` (a double atan2 result with an error of a few ulps) to float could round differently from the exact value:
// the 29 bits below the float mantissa are within 8 of the midpoint pattern, or the float result would be subnormal
HGB_HD inline bool near_float_midpoint(double d) {
	const uint64_t u = d2u(d);
	const uint32_t e = uint32_t(u >> 52) & 0x7ffu;
	if (e == 0x7ffu)
		return false;  // inf / nan: nothing to refine
	if (e < 1023u - 125u)
		return (u << 1) != 0; // below 2^-125 (float subnormals round at another bit); +-0 is exact
	const int64_t rem = int64_t(u & 0x1fffffffu);
	const int64_t dist = rem - (int64_t(1) << 28);
	return dist >= -8 && dist <= 8;
}

// (float) atan2((double) y, (double) x) of a correctly rounded libm. `lib` = the platform's double atan2 of the same arguments
// (<= 2 ulp): CUDA's atan2 on the device, anything of that accuracy on the host.
HGB_HD inline float atan2f_ref_from(double lib, float y, float x) {
	if (!near_float_midpoint(lib) || x == 0.f || y == 0.f || !(fabsf(x) <= 3.402823466e+38f) || !(fabsf(y) <= 3.402823466e+38f))
		return float(lib);
	const dd r = atan2_dd(double(y), double(x));
	return float(r.hi); // r is normalised: r.hi = RN_double(hi + lo), the double a correctly rounded atan2 returns
}

} // namespace hgb_exact
