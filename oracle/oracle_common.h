// TEST INFRASTRUCTURE ONLY -- CPU oracle for the instant-NGP NeRF hot path.
//
// This directory is a CPU restatement (C++/OpenMP) of the arithmetic of the
// reference's GPU-only kernels.  It is imported only by tests/, by
// __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference
// legs, always as the checker or the timed CPU baseline, never by the product.
//
// PARITY PINNED to the reference's own CUDA build: the reference ships no golden
// vectors or tests of its own and is GPU-only, so
// oracle/build_ref.sh compiles the unmodified reference for sm_100 into
// oracle/_ref/ref_harness, scripts/make_golden_from_ref.py ran it on a B200 and
// wrote tests/golden/ref_cuda_step.npz, and tests/test_oracle_vs_reference_golden.py
// checks this restatement against those vectors (integer outcomes and sample
// coordinates bit-exact).  Analytic known answers in tests/ cover the rest.
//
// Conventions shared with the CUDA kernels (documented in DESIGN.md):
//  * no implicit FMA contraction (-ffp-contract=off here, -fmad=false there);
//    every fused multiply-add is an explicit fmaf() in both implementations.
//  * fp16 rounding points are explicit: h16r() rounds a double/float to the
//    nearest fp16 (ties to even) and returns it as float.
#pragma once
#include <cstdint>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <vector>

namespace orc {

typedef _Float16 h16;

static inline float h16r(float x) { return (float)(h16)x; }
static inline float h16r_d(double x) { return (float)(h16)x; }
// fp16 + fp16 -> fp16, single rounding (the exact sum fits a double).
static inline float h16add(float a, float b) { return h16r_d((double)a + (double)b); }
static inline uint16_t h16bits(float x) { h16 h = (h16)x; uint16_t u; std::memcpy(&u, &h, 2); return u; }
static inline float h16from(uint16_t u) { h16 h; std::memcpy(&h, &u, 2); return (float)h; }

// ---------------------------------------------------------------------------
// pcg32  (reference: dependencies/tiny-cuda-nn/dependencies/pcg32/pcg32.h:44-205)
// ---------------------------------------------------------------------------
struct Pcg32 {
	uint64_t state, inc;
	static constexpr uint64_t MULT = 0x5851f42d4c957f2dULL;
	Pcg32() : state(0x853c49e6748fea9bULL), inc(0xda3e39cb94b95bdbULL) {}
	explicit Pcg32(uint64_t initstate, uint64_t initseq = 1u) { seed(initstate, initseq); }
	void seed(uint64_t initstate, uint64_t initseq = 1u) {
		state = 0u;
		inc = (initseq << 1u) | 1u;
		next_uint();
		state += initstate;
		next_uint();
	}
	uint32_t next_uint() {
		uint64_t old = state;
		state = old * MULT + inc;
		uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
		uint32_t rot = (uint32_t)(old >> 59u);
		return (xs >> rot) | (xs << ((~rot + 1u) & 31));
	}
	float next_float() {
		uint32_t u = (next_uint() >> 9) | 0x3f800000u;
		float f; std::memcpy(&f, &u, 4);
		return f - 1.0f;
	}
	void advance(int64_t delta_ = (1ll << 32)) {
		uint64_t cur_mult = MULT, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
		uint64_t delta = (uint64_t)delta_;
		while (delta > 0) {
			if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
			cur_plus = (cur_mult + 1) * cur_plus;
			cur_mult *= cur_mult;
			delta /= 2;
		}
		state = acc_mult * state + acc_plus;
	}
};

// ---------------------------------------------------------------------------
// Morton helpers (reference: tiny-cuda-nn/common_device.h:340-364)
// ---------------------------------------------------------------------------
static inline uint32_t expand_bits(uint32_t v) {
	v = (v * 0x00010001u) & 0xFF0000FFu;
	v = (v * 0x00000101u) & 0x0F00F00Fu;
	v = (v * 0x00000011u) & 0xC30C30C3u;
	v = (v * 0x00000005u) & 0x49249249u;
	return v;
}
static inline uint32_t morton3D(uint32_t x, uint32_t y, uint32_t z) {
	return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2);
}
static inline uint32_t morton3D_invert(uint32_t x) {
	x = x & 0x49249249;
	x = (x | (x >> 2)) & 0xc30c30c3;
	x = (x | (x >> 4)) & 0x0f00f00f;
	x = (x | (x >> 8)) & 0xff0000ff;
	x = (x | (x >> 16)) & 0x0000ffff;
	return x;
}

// ---------------------------------------------------------------------------
// colour space (reference: internal/utils/common_device.h:71-101)
// ---------------------------------------------------------------------------
static inline float srgb_to_linear(float s) {
	return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f);
}
static inline float linear_to_srgb(float l) {
	return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f;
}

struct V3 { float x, y, z; };

// wire formats (reference: internal/utils/common.h:114-153)
struct Coordinate { float px, py, pz, dt, dx, dy, dz; };           // 28 B
struct RayPayload { uint32_t alive; float ox, oy, oz, dx, dy, dz; uint32_t idx; float t; uint32_t n_steps; }; // 40 B

static inline uint32_t next_multiple(uint32_t v, uint32_t d) { return ((v + d - 1) / d) * d; }

} // namespace orc
