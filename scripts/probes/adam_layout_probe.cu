// Does the optimizer's state layout limit its DRAM efficiency?  Same bytes per parameter group, two layouts:
//   A: four flat fp32/u32 arrays (w32, m1, m2, steps) as the reference keeps them (13 streams with g16 / w16 / ema)
//   B: one 128-byte record per group of 8 parameters [w32 x8 | m1 x8 | m2 x8 | steps x8] (7 streams)
// plus a plain copy of the same number of bytes for scale.  Arithmetic is a stand-in (a few FMAs): the question is memory only.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o adam_layout_probe adam_layout_probe.cu ; run: ./adam_layout_probe [touched_fraction=0.8]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ float4 upd(float4 a, float4 b) { return make_float4(a.x * 0.9f + b.x, a.y * 0.9f + b.y, a.z * 0.9f + b.z, a.w * 0.9f + b.w); }

__global__ void __launch_bounds__(256, 3) flat_kernel(uint64_t n8, float4* w32, float4* m1, float4* m2, uint4* steps, const uint4* g16, uint4* w16, uint4* ema) {
	for (uint64_t q = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; q < n8; q += (uint64_t)gridDim.x * blockDim.x) {
		const uint4 g = __ldg(g16 + q); const uint4 w = w16[q]; uint4 e = ema[q];
		if ((g.x | g.y | g.z | g.w) != 0u) {
			float4 a0 = w32[2 * q], a1 = w32[2 * q + 1], b0 = m1[2 * q], b1 = m1[2 * q + 1], c0 = m2[2 * q], c1 = m2[2 * q + 1];
			uint4 s0 = steps[2 * q], s1 = steps[2 * q + 1];
			b0 = upd(b0, a0); b1 = upd(b1, a1); c0 = upd(c0, b0); c1 = upd(c1, b1); a0 = upd(a0, c0); a1 = upd(a1, c1); s0.x++; s1.x++;
			w32[2 * q] = a0; w32[2 * q + 1] = a1; m1[2 * q] = b0; m1[2 * q + 1] = b1; m2[2 * q] = c0; m2[2 * q + 1] = c1; steps[2 * q] = s0; steps[2 * q + 1] = s1;
			uint4 nw = w; nw.x ^= __float_as_uint(a0.x); w16[q] = nw;
		}
		e.x += w.x; ema[q] = e;
	}
}

__global__ void __launch_bounds__(256, 3) packed_kernel(uint64_t n8, float4* state /* 8 float4 per group */, const uint4* g16, uint4* w16, uint4* ema) {
	for (uint64_t q = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; q < n8; q += (uint64_t)gridDim.x * blockDim.x) {
		const uint4 g = __ldg(g16 + q); const uint4 w = w16[q]; uint4 e = ema[q];
		if ((g.x | g.y | g.z | g.w) != 0u) {
			float4* r = state + q * 8;
			float4 a0 = r[0], a1 = r[1], b0 = r[2], b1 = r[3], c0 = r[4], c1 = r[5], s0 = r[6], s1 = r[7];
			b0 = upd(b0, a0); b1 = upd(b1, a1); c0 = upd(c0, b0); c1 = upd(c1, b1); a0 = upd(a0, c0); a1 = upd(a1, c1); s0.x += 1.f; s1.x += 1.f;
			r[0] = a0; r[1] = a1; r[2] = b0; r[3] = b1; r[4] = c0; r[5] = c1; r[6] = s0; r[7] = s1;
			uint4 nw = w; nw.x ^= __float_as_uint(a0.x); w16[q] = nw;
		}
		e.x += w.x; ema[q] = e;
	}
}

__global__ void copy_kernel(uint64_t n, const uint4* a, uint4* b) {
	for (uint64_t q = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; q < n; q += (uint64_t)gridDim.x * blockDim.x) b[q] = a[q];
}

int main(int argc, char** argv) {
	const double frac = argc > 1 ? atof(argv[1]) : 0.8;
	const uint64_t P = 12206480 / 8 * 8, n8 = P / 8;
	float4 *w32, *m1, *m2, *state; uint4 *steps, *g16, *w16, *ema, *big_a, *big_b;
	CK(cudaMalloc(&w32, P * 4)); CK(cudaMalloc(&m1, P * 4)); CK(cudaMalloc(&m2, P * 4)); CK(cudaMalloc(&steps, P * 4)); CK(cudaMalloc(&state, P * 16));
	CK(cudaMalloc(&g16, P * 2)); CK(cudaMalloc(&w16, P * 2)); CK(cudaMalloc(&ema, P * 2));
	const uint64_t big = 176u << 20;               // 176 MB read + 176 MB written: the optimizer's traffic
	CK(cudaMalloc(&big_a, big)); CK(cudaMalloc(&big_b, big));
	CK(cudaMemset(w32, 0, P * 4)); CK(cudaMemset(m1, 0, P * 4)); CK(cudaMemset(m2, 0, P * 4)); CK(cudaMemset(steps, 0, P * 4)); CK(cudaMemset(state, 0, P * 16));
	CK(cudaMemset(w16, 0, P * 2)); CK(cudaMemset(ema, 0, P * 2)); CK(cudaMemset(big_a, 1, big));
	// gradients: a group is "touched" with probability frac
	uint4* h = (uint4*)malloc(n8 * 16);
	uint64_t s = 88172645463325252ull, touched = 0;
	for (uint64_t q = 0; q < n8; ++q) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; const bool t = (s >> 11) * (1.0 / 9007199254740992.0) < frac; h[q] = t ? make_uint4(1, 2, 3, 4) : make_uint4(0, 0, 0, 0); touched += t; }
	CK(cudaMemcpy(g16, h, n8 * 16, cudaMemcpyHostToDevice));
	cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
	const int grid = 148 * 6, reps = 20;
	const double bytes = n8 * (48.0 + 16.0) + touched * (128.0 * 2 + 16.0);
	float ms;
	for (int pass = 0; pass < 2; ++pass) {
		CK(cudaEventRecord(e0)); for (int r = 0; r < reps; ++r) flat_kernel<<<grid, 256>>>(n8, w32, m1, m2, steps, g16, w16, ema); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
		CK(cudaEventElapsedTime(&ms, e0, e1)); if (pass) printf("flat   : %.1f us  %.0f GB/s (%.0f MB)\n", ms * 1e3 / reps, bytes / (ms * 1e-3 / reps) / 1e9, bytes / 1e6);
		CK(cudaEventRecord(e0)); for (int r = 0; r < reps; ++r) packed_kernel<<<grid, 256>>>(n8, state, g16, w16, ema); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
		CK(cudaEventElapsedTime(&ms, e0, e1)); if (pass) printf("packed : %.1f us  %.0f GB/s\n", ms * 1e3 / reps, bytes / (ms * 1e-3 / reps) / 1e9);
		CK(cudaEventRecord(e0)); for (int r = 0; r < reps; ++r) copy_kernel<<<148 * 8, 256>>>(big / 16, big_a, big_b); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
		CK(cudaEventElapsedTime(&ms, e0, e1)); if (pass) printf("copy   : %.1f us  %.0f GB/s (352 MB)\n", ms * 1e3 / reps, 2.0 * big / (ms * 1e-3 / reps) / 1e9);
	}
	return 0;
}
