// Measured ceilings for the hash-grid kernels' access patterns on this GPU.  The table the encoder gathers from (24 MB at
// T = 2^19) lives in L2, so the honest denominator for "fraction of roofline" is not the HBM copy rate but what the L1TEX / L2
// path delivers for (a) coalesced 16-byte reads of an L2-resident buffer, (b) scattered gathers -- every lane its own line --
// and (c) scattered fp16 vector reductions.  bench.py runs these three probes (a few hundred microseconds each) next to the
// timed region and reports the encoder kernels against them; MEASURED_PEAKS.json has no L2 figure (SURVEY.md 8d asks for one).
#include "common.cuh"

namespace ngp {

__device__ __forceinline__ uint32_t mix32(uint32_t x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

// (a) every thread streams 16-byte words of the buffer, `passes` times; the buffer fits L2, the first pass warms it
__global__ void __launch_bounds__(256) l2_stream_kernel(const uint4* __restrict__ buf, uint64_t n16, uint32_t passes, uint32_t* __restrict__ sink) {
	uint32_t acc = 0;
	for (uint32_t p = 0; p < passes; ++p)
		for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x) {
			uint4 v;
			asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));   // .cg: L2 only
			acc ^= v.x ^ v.y ^ v.z ^ v.w;
		}
	if (acc == 0x12345678u) *sink = acc;
}

// (b) `per_thread` dependent-free gathers of BYTES (4 or 16) at hashed entry indices: every lane of a warp hits its own line
template <int BYTES>
__global__ void __launch_bounds__(256) l2_gather_kernel(const uint32_t* __restrict__ table, uint32_t n_entries_mask, uint32_t per_thread, uint32_t* __restrict__ sink) {
	const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
	uint32_t acc = 0;
	#pragma unroll 8
	for (uint32_t k = 0; k < per_thread; ++k) {
		const uint32_t e = mix32(tid * 0x9e3779b9u + k * 0x85ebca6bu) & n_entries_mask;
		if (BYTES == 4) acc ^= __ldg(table + e);
		else { const uint4 v = __ldg(reinterpret_cast<const uint4*>(table) + (e >> 2)); acc ^= v.x ^ v.w; }
	}
	if (acc == 0x12345678u) *sink = acc;
}

// (c) scattered reductions of BYTES (4: f16x2, 16: v4.f16x2) at hashed entry indices
template <int BYTES>
__global__ void __launch_bounds__(256) l2_red_kernel(uint32_t* __restrict__ table, uint32_t n_entries_mask, uint32_t per_thread) {
	const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
	#pragma unroll 8
	for (uint32_t k = 0; k < per_thread; ++k) {
		const uint32_t e = mix32(tid * 0x9e3779b9u + k * 0x85ebca6bu) & n_entries_mask;
		if (BYTES == 4) asm volatile("red.global.add.noftz.f16x2 [%0], %1;" :: "l"(table + e), "r"(0x80008000u) : "memory");
		else asm volatile("red.global.add.noftz.v4.f16x2 [%0], {%1, %1, %1, %1};" :: "l"(table + (e & ~3u)), "r"(0x80008000u) : "memory");
	}
}

} // namespace ngp

using namespace ngp;

// kind: 0 = coalesced L2 stream (returns GB/s), 1 / 2 = scattered 4- / 16-byte gathers, 3 / 4 = scattered 4- / 16-byte fp16 reductions
// (return 1e9 accesses per second).  d_table: table_bytes (a power of two, e.g. 32 MB: L2-resident) of scratch, zero-filled by the call.
// Times `repeats` launches with CUDA events on `stream` after one warm-up launch; synchronises the stream.
extern "C" int ngp_l2_probe(void* stream, int kind, void* d_table, uint64_t table_bytes, uint32_t repeats, double* result) {
	NGP_REQUIRE(d_table && result, "null pointer");
	NGP_REQUIRE(table_bytes >= (1u << 20) && (table_bytes & (table_bytes - 1)) == 0 && ((uintptr_t)d_table & 15u) == 0, "table_bytes must be a power of two >= 1 MB, 16-byte aligned");
	NGP_REQUIRE(kind >= 0 && kind <= 4 && repeats >= 1, "kind in [0,4], repeats >= 1");
	cudaStream_t s = (cudaStream_t)stream;
	if (cudaMemsetAsync(d_table, 0, table_bytes, s) != cudaSuccess) { set_error("ngp_l2_probe: memset failed"); return -2; }
	const uint32_t mask = (uint32_t)(table_bytes / 4 - 1), per_thread = 64, grid = kNumSMs * 8 * 4, passes = 8;
	uint32_t* sink = (uint32_t*)d_table;
	cudaEvent_t a, b;
	if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) { set_error("ngp_l2_probe: event create failed"); return -2; }
	for (uint32_t r = 0; r <= repeats; ++r) {
		if (r == 1) cudaEventRecord(a, s);
		switch (kind) {
			case 0: l2_stream_kernel<<<kNumSMs * 8, 256, 0, s>>>((const uint4*)d_table, table_bytes / 16, passes, sink); break;
			case 1: l2_gather_kernel<4><<<grid, 256, 0, s>>>((const uint32_t*)d_table, mask, per_thread, sink); break;
			case 2: l2_gather_kernel<16><<<grid, 256, 0, s>>>((const uint32_t*)d_table, mask, per_thread, sink); break;
			case 3: l2_red_kernel<4><<<grid, 256, 0, s>>>((uint32_t*)d_table, mask, per_thread); break;
			default: l2_red_kernel<16><<<grid, 256, 0, s>>>((uint32_t*)d_table, mask, per_thread); break;
		}
	}
	cudaEventRecord(b, s);
	float ms = 0.f;
	const bool ok = cudaEventSynchronize(b) == cudaSuccess && cudaEventElapsedTime(&ms, a, b) == cudaSuccess && cudaGetLastError() == cudaSuccess;
	cudaEventDestroy(a); cudaEventDestroy(b);
	if (!ok || ms <= 0.f) { set_error("ngp_l2_probe: timing failed: %s", cudaGetErrorString(cudaGetLastError())); return -2; }
	const double sec = (double)ms * 1e-3 / repeats;
	if (kind == 0) *result = (double)table_bytes * passes / sec / 1e9;
	else *result = (double)grid * 256.0 * per_thread / sec / 1e9;
	return 0;
}
