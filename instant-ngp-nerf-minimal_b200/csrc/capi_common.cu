// Error plumbing of the C ABI plus the two small utility kernels (dataset conversion, loss reduction).
#include "common.cuh"
#include <stdarg.h>

namespace ngp {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
	va_list ap; va_start(ap, fmt);
	vsnprintf(g_error, sizeof(g_error), fmt, ap);
	va_end(ap);
}

int check_launch(const char* what) {
	cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) { set_error("%s: CUDA launch failed: %s", what, cudaGetErrorString(e)); return -2; }
	return 0;
}

// internal/utils/common_device.h:160-175: 8-bit sRGB RGBA -> linear, alpha-premultiplied fp16
__global__ void from_rgba32_kernel(uint64_t n, const uint32_t* __restrict__ in, uint2* __restrict__ out) {
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
		const uint32_t p = in[i];
		const float a = (float)(p >> 24) * (1.0f / 255.0f);
		const float r = srgb_to_linear((float)(p & 0xff) * (1.0f / 255.0f)) * a;
		const float g = srgb_to_linear((float)((p >> 8) & 0xff) * (1.0f / 255.0f)) * a;
		const float b = srgb_to_linear((float)((p >> 16) & 0xff) * (1.0f / 255.0f)) * a;
		__half2 lo = __floats2half2_rn(r, g), hi = __floats2half2_rn(b, a);
		uint2 o; o.x = *reinterpret_cast<uint32_t*>(&lo); o.y = *reinterpret_cast<uint32_t*>(&hi);
		out[i] = o;
	}
}

// deterministic two-level tree: fixed CTA count, each CTA sums a contiguous slice, last CTA (ticket) sums the partials in order
__global__ void __launch_bounds__(256) reduce_sum_kernel(const float* __restrict__ in, uint64_t n, float* __restrict__ out, float* partials, unsigned int* ticket) {
	__shared__ float sh[256];
	__shared__ bool last;
	const uint64_t per = (n + gridDim.x - 1) / gridDim.x;
	const uint64_t lo = per * blockIdx.x, hi = lo + per < n ? lo + per : n;
	float s = 0.f;
	for (uint64_t i = lo + threadIdx.x; i < hi; i += 256) s += in[i];
	sh[threadIdx.x] = s;
	__syncthreads();
	for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
	if (threadIdx.x == 0) {
		partials[blockIdx.x] = sh[0];
		__threadfence();
		last = atomicAdd(ticket, 1u) == gridDim.x - 1;
	}
	__syncthreads();
	if (last) {
		__threadfence();
		float t = 0.f;
		for (uint32_t i = threadIdx.x; i < gridDim.x; i += 256) t += ((volatile float*)partials)[i];
		sh[threadIdx.x] = t;
		__syncthreads();
		for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
		if (threadIdx.x == 0) { *out = sh[0]; *ticket = 0; }
	}
}

} // namespace ngp

using namespace ngp;

extern "C" const char* ngp_last_error(void) { return g_error; }
extern "C" int ngp_abi_version(void) { return 1; }

extern "C" int ngp_from_rgba32(void* stream, uint64_t n_pixels, const uint8_t* d_rgba8, void* d_out_half) {
	NGP_REQUIRE(d_rgba8 && d_out_half, "null pointer");
	if (n_pixels == 0) return 0;
	from_rgba32_kernel<<<grid_for(n_pixels, 256), 256, 0, (cudaStream_t)stream>>>(n_pixels, (const uint32_t*)d_rgba8, (uint2*)d_out_half);
	return check_launch("from_rgba32");
}

// d_out: 1 float result followed by >= 1 + 592 floats of scratch owned by the caller (partials + ticket), i.e.
// the caller passes a buffer of at least 600 floats whose ticket word (index 1) is zero on first use.
extern "C" int ngp_reduce_sum(void* stream, const float* d_in, uint64_t n, float* d_out) {
	NGP_REQUIRE(d_in && d_out, "null pointer");
	const uint32_t grid = grid_for(n, 256, 4);
	reduce_sum_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_in, n, d_out, d_out + 2, (unsigned int*)(d_out + 1));
	return check_launch("reduce_sum");
}
