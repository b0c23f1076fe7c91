// Error plumbing of the C ABI plus the two small utility kernels (dataset conversion, loss reduction).
#include "common.cuh"
#include <stdarg.h>
#include <stdlib.h>
#include <vector>
#include <string>
#include <map>

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

// ---- launch counter + optional per-kernel event timing -------------------------------------------------
static unsigned long long g_launches = 0;
static bool g_timing = false;
struct TimedLaunch { const char* name; cudaEvent_t a, b; bool used; };
static std::vector<TimedLaunch> g_pool;
static size_t g_pool_used = 0;

bool chained_launch_enabled() {
	return !g_timing;          // per-kernel timing wants each kernel alone between its two events
}

KernelTimer::KernelTimer(const char* name, cudaStream_t s) : slot(-1), stream(s) {
	++g_launches;
	if (!g_timing || g_pool_used >= g_pool.size()) return;
	slot = (int)g_pool_used++;
	g_pool[slot].name = name;
	g_pool[slot].used = true;
	cudaEventRecord(g_pool[slot].a, stream);
}
KernelTimer::~KernelTimer() { if (slot >= 0) cudaEventRecord(g_pool[slot].b, stream); }

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

extern "C" unsigned long long ngp_launch_count(void) { return g_launches; }

extern "C" int ngp_kernel_timing_enable(int on, uint32_t max_launches) {
	if (on && g_pool.size() < max_launches) {
		const size_t old = g_pool.size();
		g_pool.resize(max_launches);
		for (size_t i = old; i < g_pool.size(); ++i) {
			g_pool[i].used = false; g_pool[i].name = "";
			if (cudaEventCreate(&g_pool[i].a) != cudaSuccess || cudaEventCreate(&g_pool[i].b) != cudaSuccess) { set_error("ngp_kernel_timing_enable: cudaEventCreate failed"); return -2; }
		}
	}
	g_timing = on != 0;
	return 0;
}

// Waits for the device, sums the recorded launches by kernel name and resets the recorder.  names: [max_kernels][64] chars.
extern "C" int ngp_kernel_timing_collect(uint32_t max_kernels, char* names, float* total_ms, uint32_t* counts) {
	if (cudaDeviceSynchronize() != cudaSuccess) { set_error("ngp_kernel_timing_collect: device synchronise failed: %s", cudaGetErrorString(cudaGetLastError())); return -2; }
	std::map<std::string, std::pair<double, uint32_t>> acc;
	for (size_t i = 0; i < g_pool_used; ++i) {
		float ms = 0.f;
		if (cudaEventElapsedTime(&ms, g_pool[i].a, g_pool[i].b) != cudaSuccess) { cudaGetLastError(); continue; }
		auto& e = acc[g_pool[i].name];
		e.first += ms; e.second += 1;
	}
	g_pool_used = 0;
	uint32_t k = 0;
	for (auto& kv : acc) {
		if (k >= max_kernels) break;
		snprintf(names + (size_t)k * 64, 64, "%s", kv.first.c_str());
		total_ms[k] = (float)kv.second.first; counts[k] = kv.second.second;
		++k;
	}
	return (int)k;
}

// Per-launch timeline of the recorded region: start / end of launch i in ms relative to the first recorded launch, in launch
// order (events on different streams of one device share a clock).  Synchronises the device and resets the recorder.
extern "C" int ngp_kernel_timing_timeline(uint32_t max_launches, char* names, float* start_ms, float* end_ms) {
	if (cudaDeviceSynchronize() != cudaSuccess) { set_error("ngp_kernel_timing_timeline: device synchronise failed: %s", cudaGetErrorString(cudaGetLastError())); return -2; }
	uint32_t k = 0;
	for (size_t i = 0; i < g_pool_used && k < max_launches; ++i) {
		float a = 0.f, b = 0.f;
		if (cudaEventElapsedTime(&a, g_pool[0].a, g_pool[i].a) != cudaSuccess || cudaEventElapsedTime(&b, g_pool[0].a, g_pool[i].b) != cudaSuccess) { cudaGetLastError(); continue; }
		snprintf(names + (size_t)k * 64, 64, "%s", g_pool[i].name);
		start_ms[k] = a; end_ms[k] = b;
		++k;
	}
	g_pool_used = 0;
	return (int)k;
}

extern "C" int ngp_from_rgba32(void* stream, uint64_t n_pixels, const uint8_t* d_rgba8, void* d_out_half) {
	NGP_REQUIRE(d_rgba8 && d_out_half, "null pointer");
	if (n_pixels == 0) return 0;
	NGP_LAUNCH("from_rgba32_kernel", (cudaStream_t)stream, from_rgba32_kernel<<<grid_for(n_pixels, 256), 256, 0, (cudaStream_t)stream>>>(n_pixels, (const uint32_t*)d_rgba8, (uint2*)d_out_half));
	return check_launch("from_rgba32");
}

// d_out: 1 float result followed by >= 1 + 592 floats of scratch owned by the caller (partials + ticket), i.e.
// the caller passes a buffer of at least 600 floats whose ticket word (index 1) is zero on first use.
extern "C" int ngp_reduce_sum(void* stream, const float* d_in, uint64_t n, float* d_out) {
	NGP_REQUIRE(d_in && d_out, "null pointer");
	const uint32_t grid = grid_for(n, 256, 4);
	NGP_LAUNCH("reduce_sum_kernel", (cudaStream_t)stream, reduce_sum_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(d_in, n, d_out, d_out + 2, (unsigned int*)(d_out + 1)));
	return check_launch("reduce_sum");
}
