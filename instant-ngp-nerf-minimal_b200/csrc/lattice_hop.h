// The empty-voxel hop of the reference's ray march in closed form.
//
// advance_to_next_voxel (internal/sampler/occupancy_sampler.cu:284-294) ends in  do { t += dt; } while (t < t_target);  with the
// base configuration's constant step (cone_angle 0: dt == min_step everywhere) that loop walks an exact arithmetic progression:
// while t stays inside one binade [2^e, 2^(e+1)) every t is a multiple of u = 2^(e-23), and fl(t + dt) = t + dtq with dtq = dt
// rounded to a multiple of u -- unless dt lies exactly half way between two multiples (a tie, broken by the parity of t).  So
//     k = max(1, ceil((t_target - t) / dtq)),   t' = t + k dtq
// is the loop's result bit for bit, as long as t' is still below the top of the binade (the sum is then exactly representable and
// the fused multiply-add returns it unrounded).  The quotient is taken in float and corrected against the exact lattice values, so
// no assumption is made about its rounding.  Every case outside these conditions reports failure and the caller runs the loop.
//
// Plain C/C++ so that the CPU test suite can compile it with gcc and compare it with the loop on millions of random inputs
// (tests/test_lattice_hop.py).
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define NGP_HD __host__ __device__ __forceinline__
#else
#define NGP_HD static inline
#endif

NGP_HD float ngp_binade_top(float t) {            // 2^(e+1) for t in [2^e, 2^(e+1)), t > 0 and normal
	uint32_t b;
	memcpy(&b, &t, 4);
	b = (b & 0x7f800000u) + 0x00800000u;
	float r;
	memcpy(&r, &b, 4);
	return r;
}

// 1 and *t_out = the loop's result, or 0 (caller must run the loop).  dt is the constant step.
NGP_HD int ngp_lattice_hop(float t, float t_target, float dt, float* t_out) {
	if (!(t >= 1.1754944e-38f) || !(dt > 0.f) || !(t_target < 3.0e38f)) return 0;
	const float top = ngp_binade_top(t);
	const float u = top * 5.9604645e-8f;          // 2^-24 * 2^(e+1) = ulp of the binade
	const float t1 = t + dt;
	if (!(t1 < top)) return 0;
	const float dtq = t1 - t;                     // exact (both multiples of u, t1 < 2 t)
	if (!(dtq > 0.f) || fabsf(dt - dtq) == 0.5f * u) return 0;      // a tie: the progression depends on the parity of t
	float k = ceilf((t_target - t) / dtq);
	if (!(k >= 1.f)) k = 1.f;
	if (!(k < 8388608.f)) return 0;
	float tc = fmaf(k, dtq, t);
	if (!(tc < top)) return 0;
	// the float quotient is off by at most one: settle on the first lattice point that is not below the target
	if (tc < t_target) {
		k += 1.f; tc = fmaf(k, dtq, t);
		if (!(tc < top) || tc < t_target) return 0;
	} else if (k > 1.f) {
		const float tp = fmaf(k - 1.f, dtq, t);
		if (!(tp < t_target)) {
			k -= 1.f; tc = tp;
			if (k > 1.f && !(fmaf(k - 1.f, dtq, t) < t_target)) return 0;
		}
	}
	*t_out = tc;
	return 1;
}

// The same in two halves for a marching loop that makes many hops inside one binade: the per-binade constants are built once
// (ngp_lattice_enter) and a hop is then a multiplication by the cached reciprocal, a ceil and the fix-up against exact lattice
// values.  The constants stay valid while t is below `top`; the caller re-enters after any step that takes t to `top` or beyond,
// and after every hop the closed form declined (which it then made with the loop).
typedef struct { float top, dtq, inv_dtq; } ngp_lattice;     /* inv_dtq == 0: no closed form in this binade */

NGP_HD void ngp_lattice_enter(ngp_lattice* L, float t, float dt, int const_dt) {
	L->top = 0.f; L->dtq = 0.f; L->inv_dtq = 0.f;
	if (!const_dt || !(t >= 1.1754944e-38f) || !(dt > 0.f) || !(t < 1.0e38f)) return;
	const float top = ngp_binade_top(t), u = top * 5.9604645e-8f;
	L->top = top;
	const float t1 = t + dt;
	if (!(t1 < top)) return;                                   /* the very next point leaves the binade: step by hand */
	const float dtq = t1 - t;
	if (!(dtq > 0.f) || fabsf(dt - dtq) == 0.5f * u) return;   /* a tie: the progression depends on the parity of t */
	L->dtq = dtq; L->inv_dtq = 1.0f / dtq;
}

NGP_HD int ngp_lattice_hop_cached(const ngp_lattice* L, float t, float t_target, float* t_out) {
	if (L->inv_dtq == 0.f || !(t_target < 3.0e38f) || !(t < L->top)) return 0;
	float k = fmaxf(ceilf((t_target - t) * L->inv_dtq), 1.f);
	if (!(k < 8388608.f)) return 0;
	float tc = fmaf(k, L->dtq, t);
	if (!(tc < L->top)) return 0;
	/* the approximate quotient is off by at most one: settle on the first lattice point that is not below the target */
	if (tc < t_target) {
		k += 1.f; tc = fmaf(k, L->dtq, t);
		if (!(tc < L->top) || tc < t_target) return 0;
	} else if (k > 1.f) {
		const float tp = fmaf(k - 1.f, L->dtq, t);
		if (!(tp < t_target)) {
			if (k > 2.f && !(fmaf(k - 2.f, L->dtq, t) < t_target)) return 0;
			tc = tp;
		}
	}
	*t_out = tc;
	return 1;
}
