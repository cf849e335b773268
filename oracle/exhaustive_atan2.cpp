// TEST INFRASTRUCTURE. Host check of harfang3d_b200/csrc/hgb_atan2_exact.h — the product's restatement of
// "(float) atan2((double) y, (double) x)" as the reference's ToEuler computes it (foundation/matrix3.cpp:91-137) — against the libm the
// reference links. 2^64 argument pairs cannot be enumerated; this samples N pairs per thread from three distributions (raw bit
// patterns, rotation-matrix-like values in [-1, 1], ratios close to 1), and for each pair plays the role of a 2-ulp device library by
// perturbing libm's double by -2 ... +2 ulps before handing it to atan2f_ref_from(): the result must equal the narrowed libm value
// every time. The double-double core is additionally compared with libquadmath on a subset (relative error bound 2^-95), and the
// number of samples that really took the slow path is reported.
//   g++ -O2 -mfma -ffp-contract=off -I../harfang3d_b200/csrc exhaustive_atan2.cpp -lquadmath -lpthread
#include "hgb_atan2_exact.h"
#include <quadmath.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>

static inline uint64_t rng_next(uint64_t *s) { // xorshift64*
	uint64_t x = *s;
	x ^= x >> 12, x ^= x << 25, x ^= x >> 27;
	*s = x;
	return x * 0x2545F4914F6CDD1DULL;
}
static inline float u2f(uint32_t u) {
	float f;
	memcpy(&f, &u, 4);
	return f;
}
struct Job {
	uint64_t seed, n, mismatches, slow, quad_bad, forced;
	double worst_rel;
};

static void *run(void *arg) {
	Job *j = (Job *)arg;
	uint64_t s = j->seed;
	for (uint64_t i = 0; i < j->n; ++i) {
		const uint64_t r = rng_next(&s), r2 = rng_next(&s);
		float y, x;
		switch (r2 >> 62) {
		case 0: // raw bit patterns (every exponent, both signs)
			y = u2f(uint32_t(r)), x = u2f(uint32_t(r >> 32));
			break;
		case 1: // rotation-matrix-like: uniform in [-1, 1]
			y = float(int32_t(uint32_t(r))) * 4.656612873077393e-10f, x = float(int32_t(uint32_t(r >> 32))) * 4.656612873077393e-10f;
			break;
		case 2: { // same exponent, random mantissas: angles near +-pi/4, +-3pi/4
			const uint32_t e = 100u + uint32_t(r2 & 63u);
			y = u2f((uint32_t(r) & 0x807fffffu) | (e << 23)), x = u2f((uint32_t(r >> 32) & 0x807fffffu) | (e << 23));
			break;
		}
		default: // small angle / near the axes: one component tiny
			y = float(int32_t(uint32_t(r))) * 4.656612873077393e-10f * u2f(((uint32_t(r2) & 63u) + 64u) << 23), x = float(int32_t(uint32_t(r >> 32))) * 4.656612873077393e-10f;
			break;
		}
		if (y != y || x != x)
			continue;
		const double lib = atan2(double(y), double(x));
		const float want = float(lib);
		// a 2-ulp device library
		const int k = int((r2 >> 8) % 5) - 2;
		double pert = lib;
		for (int q = 0; q < (k < 0 ? -k : k); ++q)
			pert = nextafter(pert, k < 0 ? -INFINITY : INFINITY);
		if (lib == 0.0 || isinf(lib))
			pert = lib;
		const float got = hgb_exact::atan2f_ref_from(pert, y, x);
		if (memcmp(&got, &want, 4) != 0) {
			if (j->mismatches++ < 5)
				fprintf(stderr, "MISMATCH y=%a x=%a lib=%a pert=%a want=%a got=%a\n", y, x, lib, pert, want, got);
		}
		if (hgb_exact::near_float_midpoint(pert) && x != 0.f && y != 0.f && fabsf(x) <= 3.402823466e+38f && fabsf(y) <= 3.402823466e+38f)
			++j->slow;
		if ((i & 63) == 0 && x != 0.f && y != 0.f && fabsf(x) <= 3.402823466e+38f && fabsf(y) <= 3.402823466e+38f) { // the double-double core on everything, not only near midpoints
			++j->forced;
			const hgb_exact::dd t = hgb_exact::atan2_dd(double(y), double(x));
			const float f2 = float(t.hi);
			if (memcmp(&f2, &want, 4) != 0 && j->mismatches++ < 5)
				fprintf(stderr, "DD MISMATCH y=%a x=%a lib=%a dd.hi=%a\n", y, x, lib, t.hi);
			const __float128 ref = atan2q((__float128)y, (__float128)x);
			const __float128 err = fabsq(((__float128)t.hi + (__float128)t.lo) - ref);
			const double rel = double(err / fabsq(ref));
			if (rel > j->worst_rel)
				j->worst_rel = rel;
			if (rel > 0x1p-95)
				++j->quad_bad;
		}
	}
	return NULL;
}

int main(int argc, char **argv) {
	const uint64_t per_thread = argc > 1 ? strtoull(argv[1], NULL, 10) : 125000000ULL;
	const int n_thr = argc > 2 ? atoi(argv[2]) : 8;
	pthread_t th[64];
	Job jobs[64];
	for (int t = 0; t < n_thr; ++t) {
		jobs[t] = Job{0x9E3779B97F4A7C15ULL * uint64_t(t + 1), per_thread, 0, 0, 0, 0, 0.0};
		pthread_create(&th[t], NULL, run, &jobs[t]);
	}
	uint64_t mism = 0, slow = 0, qbad = 0, forced = 0;
	double worst = 0;
	for (int t = 0; t < n_thr; ++t) {
		pthread_join(th[t], NULL);
		mism += jobs[t].mismatches, slow += jobs[t].slow, qbad += jobs[t].quad_bad, forced += jobs[t].forced;
		if (jobs[t].worst_rel > worst)
			worst = jobs[t].worst_rel;
	}
	printf("samples %llu, mismatches %llu, slow-path (near a float midpoint) %llu, double-double core forced on %llu (worst rel. error vs libquadmath %.3g = 2^%.1f, above 2^-95: %llu)\n",
		(unsigned long long)(per_thread * n_thr), (unsigned long long)mism, (unsigned long long)slow, (unsigned long long)forced, worst, log2(worst > 0 ? worst : 1e-300),
		(unsigned long long)qbad);
	return mism || qbad ? 1 : 0;
}
