// Shared device/host helpers for the sm_100a kernels.
// Compiled with -fmad=false: every fused multiply-add is an explicit __fmaf_rn / fmaf so integer-valued
// results (hash indices, occupancy decisions, per-ray sample counts) are reproducible against oracle/.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/ngp_b200.h"

namespace ngp {

// ---- error plumbing (C ABI returns int, message kept thread-local) ------------------------------
void set_error(const char* fmt, ...);
int check_launch(const char* what);

// Every kernel launch goes through NGP_LAUNCH: it counts the launch (ngp_launch_count) and, when kernel timing is enabled
// (ngp_kernel_timing_enable), brackets it with a pair of CUDA events on the launching stream so bench.py can report live
// per-kernel durations for the roofline without a profiler attached.
struct KernelTimer {
	KernelTimer(const char* name, cudaStream_t stream);
	~KernelTimer();
	int slot;
	cudaStream_t stream;
};
#define NGP_LAUNCH(name, stream, ...) do { ngp::KernelTimer _kt(name, stream); __VA_ARGS__; } while (0)

// Programmatic dependent launch (sm_90+).  The kernels of one training step form a dependent chain on one stream; launched
// this way, the next kernel's grid is set up and its CTAs become resident while the previous kernel drains, instead of
// after it has retired (3-4 us per boundary, ~12 boundaries per 0.6 ms step).  Every kernel launched through launch_chained
// calls chain_enter() as its FIRST statement: it blocks until the preceding grid has completed and its writes are visible,
// then lets its own successor start launching.  Waiting before triggering keeps at most two grids in flight, and because a
// grid only triggers after all its CTAs have passed the wait, completion is transitive along the chain.
bool chained_launch_enabled();     // off while per-kernel timing is enabled (each kernel alone between its two events)
template <typename... Exp, typename... Act>
inline cudaError_t launch_chained(void (*kernel)(Exp...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Act&&... args) {
	cudaLaunchConfig_t cfg = {};
	cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
	cudaLaunchAttribute attr[1];
	attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
	attr[0].val.programmaticStreamSerializationAllowed = 1;
	cfg.attrs = attr; cfg.numAttrs = chained_launch_enabled() ? 1 : 0;
	return cudaLaunchKernelEx(&cfg, kernel, static_cast<Exp>(args)...);
}
#ifdef __CUDACC__
__device__ __forceinline__ void chain_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void chain_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void chain_enter() { chain_wait(); chain_trigger(); }
#endif

#define NGP_REQUIRE(cond, msg) do { if (!(cond)) { ngp::set_error("%s: %s", __func__, msg); return -1; } } while (0)

constexpr uint32_t kNumSMs = 148;

static inline uint32_t div_up(uint64_t a, uint32_t b) { return (uint32_t)((a + b - 1) / b); }
// grid for a grid-stride kernel: enough CTAs to fill the machine, never more than the work
static inline uint32_t grid_for(uint64_t n, uint32_t block, uint32_t ctas_per_sm = 8) {
	uint64_t need = (n + block - 1) / block;
	uint64_t cap = (uint64_t)kNumSMs * ctas_per_sm;
	return (uint32_t)(need < cap ? (need ? need : 1) : cap);
}

// ---- pcg32 (tiny-cuda-nn/dependencies/pcg32/pcg32.h:44-205) ---------------------------------------
struct Pcg32 {
	uint64_t state, inc;
	__host__ __device__ uint32_t next_uint() {
		uint64_t old = state;
		state = old * 0x5851f42d4c957f2dULL + inc;
		uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
		uint32_t rot = (uint32_t)(old >> 59u);
		return (xs >> rot) | (xs << ((~rot + 1u) & 31));
	}
	__host__ __device__ float next_float() {
		uint32_t u = (next_uint() >> 9) | 0x3f800000u;
#ifdef __CUDA_ARCH__
		return __uint_as_float(u) - 1.0f;
#else
		float f; memcpy(&f, &u, 4); return f - 1.0f;
#endif
	}
	__host__ __device__ void advance(uint64_t delta) {
		uint64_t cur_mult = 0x5851f42d4c957f2dULL, cur_plus = inc, acc_mult = 1u, acc_plus = 0u;
		while (delta > 0) {
			if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
			cur_plus = (cur_mult + 1) * cur_plus;
			cur_mult *= cur_mult;
			delta >>= 1;
		}
		state = acc_mult * state + acc_plus;
	}
	__host__ __device__ void seed(uint64_t initstate, uint64_t initseq = 1u) {
		state = 0u; inc = (initseq << 1u) | 1u; next_uint(); state += initstate; next_uint();
	}
};

// ---- Morton (tiny-cuda-nn/common_device.h:340-364) ------------------------------------------------
__host__ __device__ inline uint32_t expand_bits(uint32_t v) {
	v = (v * 0x00010001u) & 0xFF0000FFu;
	v = (v * 0x00000101u) & 0x0F00F00Fu;
	v = (v * 0x00000011u) & 0xC30C30C3u;
	v = (v * 0x00000005u) & 0x49249249u;
	return v;
}
__host__ __device__ inline uint32_t morton3D(uint32_t x, uint32_t y, uint32_t z) {
	return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2);
}
__host__ __device__ inline uint32_t morton3D_invert(uint32_t x) {
	x = x & 0x49249249;
	x = (x | (x >> 2)) & 0xc30c30c3;
	x = (x | (x >> 4)) & 0x0f00f00f;
	x = (x | (x >> 8)) & 0xff0000ff;
	x = (x | (x >> 16)) & 0x0000ffff;
	return x;
}

// ---- colour space (internal/utils/common_device.h:71-101) -----------------------------------------
__device__ inline float srgb_to_linear(float s) { return s <= 0.04045f ? s / 12.92f : powf((s + 0.055f) / 1.055f, 2.4f); }
__device__ inline float linear_to_srgb(float l) { return l < 0.0031308f ? 12.92f * l : 1.055f * powf(l, 0.41666f) - 0.055f; }

// ---- wire formats (internal/utils/common.h:114-153) -----------------------------------------------
struct Coordinate { float px, py, pz, dt, dx, dy, dz; };                     // 28 B
struct RayPayload { uint32_t alive; float ox, oy, oz, dx, dy, dz; uint32_t idx; float t; uint32_t n_steps; };  // 40 B
static_assert(sizeof(Coordinate) == 28, "Coordinate wire format");
static_assert(sizeof(RayPayload) == 40, "RayPayload wire format");

// ---- single-pass chained scan with decoupled look-back ------------------------------------------
// Workspace: [0] ticket counter, [1..] one 64-bit status word per CTA: top 2 bits flag (1 = aggregate, 2 = inclusive),
// low 62 bits value.  Values are packed (hi: count, lo: sum) by the callers and stay far below 2^62.
struct ScanState { unsigned long long* words; };

__device__ inline uint32_t scan_ticket(ScanState s) {
	__shared__ uint32_t ticket;
	if (threadIdx.x == 0) ticket = (uint32_t)atomicAdd(s.words, 1ull);
	__syncthreads();
	return ticket;
}

// Block-wide exclusive scan of one 64-bit value per thread (blockDim.x <= 1024, multiple of 32), chained across
// CTAs in ticket order.  Returns the exclusive prefix over all earlier threads of all earlier tickets.
__device__ inline unsigned long long chained_exclusive_scan(ScanState s, uint32_t ticket, unsigned long long v) {
	__shared__ unsigned long long warp_sums[32];
	__shared__ unsigned long long cta_prefix;
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
	unsigned long long incl = v;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) {
		unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
		if (lane >= o) incl += t;
	}
	if (lane == 31) warp_sums[warp] = incl;
	__syncthreads();
	if (warp == 0) {
		unsigned long long w = lane < n_warps ? warp_sums[lane] : 0ull;
		unsigned long long wi = w;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) {
			unsigned long long t = __shfl_up_sync(0xffffffffu, wi, o);
			if (lane >= o) wi += t;
		}
		warp_sums[lane] = wi - w;                      // exclusive warp offsets
		unsigned long long total = __shfl_sync(0xffffffffu, wi, 31);
		if (lane == 0) {
			volatile unsigned long long* words = s.words + 1;
			unsigned long long prefix = 0;
			if (ticket > 0) {
				words[ticket] = (1ull << 62) | total;    // publish aggregate
				__threadfence();
				int64_t j = (int64_t)ticket - 1;
				while (j >= 0) {
					unsigned long long w2;
					do { w2 = words[j]; } while ((w2 >> 62) == 0);
					prefix += w2 & ((1ull << 62) - 1);
					if ((w2 >> 62) == 2) break;
					--j;
				}
			}
			__threadfence();
			words[ticket] = (2ull << 62) | (prefix + total);  // publish inclusive prefix
			cta_prefix = prefix;
		}
	}
	__syncthreads();
	return cta_prefix + warp_sums[warp] + (incl - v);
}

} // namespace ngp
