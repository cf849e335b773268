/* TEST INFRASTRUCTURE ONLY. Compares hgp_sinf/hgp_cosf (hg_port.c) with the host libm's sinf/cosf/sincosf for ALL 2^32
 * float bit patterns. Build+run: make -C oracle exhaustive   (8 threads, ~15 s). Result recorded in DESIGN.md:
 * glibc 2.39, FMA-capable x86-64 -> 0 mismatches. */
#define _GNU_SOURCE
#include "hg_port.h"
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#define NT 8
static unsigned long long bad[NT][3];
static void *work(void *a) {
	const long id = (long)a;
	for (unsigned long long u = id; u < (1ULL << 32); u += NT) {
		unsigned ui = (unsigned)u;
		float f, s, c, ms, mc;
		memcpy(&f, &ui, 4);
		if (f != f || isinf(f))
			continue;
		ms = hgp_sinf(f), mc = hgp_cosf(f);
		s = sinf(f), c = cosf(f);
		bad[id][0] += memcmp(&s, &ms, 4) != 0;
		bad[id][1] += memcmp(&c, &mc, 4) != 0;
		sincosf(f, &s, &c); /* what g++ -O2 actually emits for matrix3.cpp:199-204 */
		bad[id][2] += memcmp(&s, &ms, 4) != 0 || memcmp(&c, &mc, 4) != 0;
	}
	return 0;
}
int main(void) {
	pthread_t t[NT];
	unsigned long long s = 0, c = 0, sc = 0;
	for (long i = 0; i < NT; ++i)
		pthread_create(&t[i], 0, work, (void *)i);
	for (int i = 0; i < NT; ++i) {
		pthread_join(t[i], 0);
		s += bad[i][0], c += bad[i][1], sc += bad[i][2];
	}
	printf("mismatches vs libm over all finite floats: sinf %llu cosf %llu sincosf %llu\n", s, c, sc);
	return (s || c || sc) ? 1 : 0;
}
