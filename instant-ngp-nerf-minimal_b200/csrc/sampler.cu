// Occupancy grid maintenance and training-ray sample generation for sm_100a.
// Replaces internal/sampler/occupancy_sampler.cu kernels K1-K8: mark_untrained_density_grid (:18),
// generate_grid_samples_nerf_nonuniform (:63), splat (:107), ema (:134), reduce_sum mean (:741), grid_to_bitfield (:160),
// bitfield_max_pool (:182), generate_training_samples_nerf (:310).
#include "march.cuh"

namespace ngp {

constexpr uint32_t kCellMul = 56924617u, kCellAdd = 96925573u;     // the cell hash of occupancy_sampler.cu:81

__global__ void mark_untrained_grid_kernel(ngp_sampler_cfg c, ngp_camera_cfg cam, uint32_t n_images, const float* __restrict__ xforms, float* __restrict__ grid_out) {
	const uint32_t G = c.grid_size, cells = G * G * G, total = cells * c.n_cascades;
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
		const uint32_t level = i / cells, pos_idx = i % cells;
		const uint32_t x = morton3D_invert(pos_idx >> 0), y = morton3D_invert(pos_idx >> 1), z = morton3D_invert(pos_idx >> 2);
		const float s = scalbnf(1.0f, (int)level), g = (float)G;
		const float px = (((float)x + 0.5f) / g - 0.5f) * s + 0.5f;
		const float py = (((float)y + 0.5f) / g - 0.5f) * s + 0.5f;
		const float pz = (((float)z + 0.5f) / g - 0.5f) * s + 0.5f;
		const float half_resx = cam.width * 0.5f, half_resy = cam.height * 0.5f;
		const float voxel_radius = 0.5f * 1.73205080757f * s / g;
		bool seen = false;
		for (uint32_t j = 0; j < n_images && !seen; ++j) {
			const float* xf = xforms + (size_t)j * 12;
			const float lx = px - xf[9], ly = py - xf[10], lz = pz - xf[11];
			// Eigen dot(): p0 + (p1 + p2), both sums contracted by nvcc
			const float cx = __fmaf_rn(lx, xf[0], __fmaf_rn(ly, xf[1], lz * xf[2]));
			const float cy = __fmaf_rn(lx, xf[3], __fmaf_rn(ly, xf[4], lz * xf[5]));
			const float cz = __fmaf_rn(lx, xf[6], __fmaf_rn(ly, xf[7], lz * xf[8]));
			if (cz > 0.f && fabsf(cx) - voxel_radius < cz / cam.fx * half_resx && fabsf(cy) - voxel_radius < cz / cam.fy * half_resy) seen = true;
		}
		grid_out[i] = seen ? 0.f : -1.f;
	}
}

// pcg32 skip-ahead through a table: entry k holds the affine map of 2^k steps (state -> state * mult + plus), so advancing by delta
// is one 64-bit multiply-add per SET bit of delta instead of the two to four multiplies per bit of Pcg32::advance's square-and-
// multiply loop (350 of the ~1 200 instructions a refresh sample cost, ncu).  Powers of one affine map commute: the state reached
// is the same 64-bit integer.
struct PcgJump { uint64_t mult[32], plus[32]; };
__device__ __forceinline__ void pcg_jump_table_build(PcgJump& t, uint64_t inc) {      // one thread
	uint64_t m = 0x5851f42d4c957f2dULL, p = inc;
	for (int k = 0; k < 32; ++k) { t.mult[k] = m; t.plus[k] = p; p = (m + 1) * p; m *= m; }
}
__device__ __forceinline__ void pcg_jump(Pcg32& rng, const PcgJump& t, uint32_t delta) {
	uint64_t st = rng.state;
	while (delta) { const int k = __ffs((int)delta) - 1; delta &= delta - 1u; st = st * t.mult[k] + t.plus[k]; }
	rng.state = st;
}

// sample i of generate_grid_samples_nerf_nonuniform (occupancy_sampler.cu:63-105): cascade from the first random number, up to 10
// pseudo-random cells until one is above the threshold, a uniform position inside the cell that was kept.  jump (optional): the
// skip-ahead table of rng0's stream; i * 4 must then fit 32 bits.
__device__ __forceinline__ void grid_sample_one(const ngp_sampler_cfg& c, uint32_t i, uint32_t n, const Pcg32& rng0, uint32_t step, const float* __restrict__ grid_in,
                                                float thresh, float* __restrict__ pos_out, uint32_t* __restrict__ idx_out, const PcgJump* jump = nullptr) {
	const uint32_t G = c.grid_size, cells = G * G * G;
	const bool cells_pow2 = (cells & (cells - 1u)) == 0u;
	Pcg32 rng = rng0;
	if (jump) pcg_jump(rng, *jump, i * 4u); else rng.advance((uint64_t)i * 4);
	const uint32_t level = (uint32_t)(rng.next_float() * (float)c.n_cascades) % c.n_cascades;
	uint32_t pos_idx = 0;
	for (uint32_t j = 0; j < 10; ++j) {
		const uint32_t h = (i + step * n) * kCellMul + j * 19349663u + kCellAdd;
		pos_idx = cells_pow2 ? (h & (cells - 1u)) : (h % cells);
		if (grid_in[pos_idx + level * cells] > thresh) break;
	}
	const uint32_t idx = pos_idx + level * cells;
	const uint32_t x = morton3D_invert(pos_idx >> 0), y = morton3D_invert(pos_idx >> 1), z = morton3D_invert(pos_idx >> 2);
	const float rx = rng.next_float(), ry = rng.next_float(), rz = rng.next_float();
	const float s = scalbnf(1.0f, (int)level), g = (float)G, diag = c.aabb_max - c.aabb_min;
	// divisions by powers of two as multiplications by their reciprocals: the same float, bit for bit
	int e;
	const bool g_pow2 = (G & (G - 1u)) == 0u, diag_pow2 = frexpf(diag, &e) == 0.5f;
	const float ig = 1.0f / g, idiag = 1.0f / diag;
	const float qx = g_pow2 ? ((float)x + rx) * ig : ((float)x + rx) / g;
	const float qy = g_pow2 ? ((float)y + ry) * ig : ((float)y + ry) / g;
	const float qz = g_pow2 ? ((float)z + rz) * ig : ((float)z + rz) / g;
	const float px = (qx - 0.5f) * s + 0.5f, py = (qy - 0.5f) * s + 0.5f, pz = (qz - 0.5f) * s + 0.5f;
	pos_out[0] = diag_pow2 ? (px - c.aabb_min) * idiag : (px - c.aabb_min) / diag;
	pos_out[1] = diag_pow2 ? (py - c.aabb_min) * idiag : (py - c.aabb_min) / diag;
	pos_out[2] = diag_pow2 ? (pz - c.aabb_min) * idiag : (pz - c.aabb_min) / diag;
	*idx_out = idx;
}

__global__ void generate_grid_samples_kernel(ngp_sampler_cfg c, uint32_t n, uint32_t i_begin, uint32_t i_end, Pcg32 rng0, uint32_t step, const float* __restrict__ grid_in, float thresh,
                                             float* __restrict__ out_pos, uint32_t* __restrict__ out_idx) {
	for (uint32_t i = i_begin + blockIdx.x * blockDim.x + threadIdx.x; i < i_end; i += gridDim.x * blockDim.x)
		grid_sample_one(c, i, n, rng0, step, grid_in, thresh, out_pos + (size_t)i * 3, out_idx + i);
}

// sum_{i<n} floor((a i + b) / m), all unsigned, in O(log m) (the Euclid-like descent)
__host__ __device__ inline uint64_t floor_sum(uint64_t n, uint64_t m, uint64_t a, uint64_t b) {
	uint64_t ans = 0;
	while (true) {
		if (a >= m) { ans += n * (n - 1) / 2 * (a / m); a %= m; }
		if (b >= m) { ans += n * (b / m); b %= m; }
		const uint64_t y_max = a * n + b;
		if (y_max < m) break;
		n = y_max / m; b = y_max % m;
		const uint64_t t = m; m = a; a = t;
	}
	return ans;
}
// how many of x_k = (alpha k + beta) mod C, k < M, are below R   (0 <= beta < C, 0 < R < C)
__host__ __device__ inline uint64_t count_below(uint64_t M, uint64_t C, uint64_t alpha, uint64_t beta, uint64_t R) {
	return M + floor_sum(M, C, alpha, beta) - floor_sum(M, C, alpha, beta + C - R);
}

// The same samples, emitted in the order of their FIRST candidate cell.  The cell hash (i + step n) * kCellMul + kCellAdd is a
// bijection modulo the power-of-two cell count, so thread m inverts it: the samples whose first candidate is Morton cell m are
// i = i0(m) + r * cells < n with i0(m) = (alpha m + beta) mod cells.  Output slots are the running count of those samples over
// the cells before m -- closed form through floor_sum, so there is no counting pass and the order is deterministic.  A warp then
// encodes 32 Morton-adjacent cells instead of 32 cells spread over the whole volume: the hash-grid gathers of the coarse
// levels hit L1 instead of L2.  The multiset of (position, index) pairs equals generate_grid_samples_kernel's.
__global__ void __launch_bounds__(256) generate_grid_samples_sorted_kernel(ngp_sampler_cfg c, uint32_t n, Pcg32 rng0, uint32_t step, uint32_t alpha, uint32_t beta, uint32_t chunk,
                                                                            uint32_t first_block, const float* __restrict__ grid_in, float thresh, float* __restrict__ out_pos, uint32_t* __restrict__ out_idx) {
	const uint32_t cells = c.grid_size * c.grid_size * c.grid_size, mask = cells - 1;
	const uint32_t K = n / cells, R = n % cells;
	const uint32_t m_begin = (first_block + blockIdx.x) * chunk, m_end = min(m_begin + chunk, cells);      // chunk and cells are multiples of 256
	if (m_begin >= cells) return;
	__shared__ uint32_t warp_tot[8];
	__shared__ uint64_t fs[2];
	__shared__ PcgJump jump;
	// one closed-form prefix per CTA (the 64-bit divisions of floor_sum are slow: two lanes of different warps take one sum each)
	if (R && (threadIdx.x == 0 || threadIdx.x == 32)) fs[threadIdx.x >> 5] = floor_sum(m_begin, cells, alpha, threadIdx.x ? (uint64_t)beta + cells - R : (uint64_t)beta);
	if (threadIdx.x == 64) pcg_jump_table_build(jump, rng0.inc);
	const bool use_jump = n <= (1u << 30);                                              // i * 4 fits 32 bits
	__syncthreads();
	uint64_t base = (uint64_t)m_begin * K + (R ? m_begin + fs[0] - fs[1] : 0);
	for (uint32_t m0 = m_begin; m0 < m_end; m0 += 256) {
		const uint32_t i0 = (alpha * (m0 + threadIdx.x) + beta) & mask;
		const uint32_t cnt = K + (i0 < R ? 1u : 0u);
		uint32_t incl = cnt;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if ((threadIdx.x & 31) >= o) incl += v; }
		__syncthreads();                                                                     // the previous round's warp_tot has been read
		if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
		__syncthreads();
		uint32_t before = 0, total = 0;
		#pragma unroll
		for (uint32_t w = 0; w < 8; ++w) { const uint32_t v = warp_tot[w]; total += v; if (w < (threadIdx.x >> 5)) before += v; }
		const uint64_t slot0 = base + before + incl - cnt;
		for (uint32_t r = 0; r < cnt; ++r)
			grid_sample_one(c, i0 + r * cells, n, rng0, step, grid_in, thresh, out_pos + (slot0 + r) * 3, out_idx + slot0 + r, use_jump ? &jump : nullptr);
		base += total;
	}
}

__device__ __forceinline__ float density_act(float v, int act) {
	switch (act) {
		case 0: return v;
		case 1: return v > 0.f ? v : 0.f;
		case 2: return 1.0f / (1.0f + expf(-v));
		default: return __expf(v);
	}
}

__global__ void splat_grid_kernel(uint32_t n, const uint32_t* __restrict__ indices, const __half* __restrict__ density, uint32_t stride, int act, float min_step, float* grid_current) {
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		const float thickness = density_act(__half2float(density[(size_t)i * stride]), act) * min_step;
		atomicMax((uint32_t*)&grid_current[indices[i]], __float_as_uint(thickness));
	}
}

__global__ void ema_grid_kernel(uint32_t n, const float* __restrict__ cur, float decay, float* __restrict__ grid) {
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		const float prev = grid[i];
		grid[i] = prev < 0.f ? prev : fmaxf(prev * decay, cur[i]);
	}
}

// mean over cascade 0 of max(v,0)/cells in two fixed-shape stages (128 CTAs of contiguous slices, then one tree over the
// partials): run-to-run deterministic, unlike the reference's float-atomic reduce_sum.
constexpr uint32_t kMeanParts = 128;
__global__ void __launch_bounds__(256) grid_mean_partial_kernel(const float* __restrict__ grid, uint32_t cells, float* __restrict__ partials) {
	__shared__ float sh[256];
	const uint32_t per = (cells + kMeanParts - 1) / kMeanParts, lo = blockIdx.x * per, hi = min(lo + per, cells);
	const float inv = (float)cells;
	float s = 0.f;
	for (uint32_t i = lo + threadIdx.x; i < hi; i += 256) s += fmaxf(grid[i], 0.f) / inv;
	sh[threadIdx.x] = s;
	__syncthreads();
	for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
	if (threadIdx.x == 0) partials[blockIdx.x] = sh[0];
}
__global__ void __launch_bounds__(kMeanParts) grid_mean_final_kernel(const float* __restrict__ partials, float* __restrict__ mean_out) {
	__shared__ float sh[kMeanParts];
	sh[threadIdx.x] = partials[threadIdx.x];
	__syncthreads();
	for (int o = kMeanParts / 2; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
	if (threadIdx.x == 0) *mean_out = sh[0];
}

__global__ void grid_to_bitfield_kernel(uint32_t n_bytes, const float* __restrict__ grid, uint8_t* __restrict__ bits, const float* __restrict__ mean, float min_optical_thickness) {
	const float thresh = fminf(min_optical_thickness, *mean);
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_bytes; i += gridDim.x * blockDim.x) {
		const float4 a = reinterpret_cast<const float4*>(grid)[(size_t)i * 2], b = reinterpret_cast<const float4*>(grid)[(size_t)i * 2 + 1];
		uint32_t v = 0;
		v |= a.x > thresh ? 1u : 0u; v |= a.y > thresh ? 2u : 0u; v |= a.z > thresh ? 4u : 0u; v |= a.w > thresh ? 8u : 0u;
		v |= b.x > thresh ? 16u : 0u; v |= b.y > thresh ? 32u : 0u; v |= b.z > thresh ? 64u : 0u; v |= b.w > thresh ? 128u : 0u;
		bits[i] = (uint8_t)v;
	}
}

__global__ void bitfield_max_pool_kernel(uint32_t n, uint32_t G, const uint8_t* __restrict__ prev, uint8_t* __restrict__ next) {
	for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
		const uint2 p = reinterpret_cast<const uint2*>(prev)[i];
		uint32_t v = 0;
		#pragma unroll
		for (int j = 0; j < 4; ++j) { v |= ((p.x >> (8 * j)) & 0xffu) ? (1u << j) : 0u; v |= ((p.y >> (8 * j)) & 0xffu) ? (16u << j) : 0u; }
		const uint32_t x = morton3D_invert(i >> 0) + G / 8, y = morton3D_invert(i >> 1) + G / 8, z = morton3D_invert(i >> 2) + G / 8;
		next[morton3D(x, y, z)] |= (uint8_t)v;
	}
}

// one thread per output byte = 8 blocks of 512 cells = 8 x 64 bytes of the cell bitfield
__global__ void coarse_bitfield_kernel(uint32_t n_bytes, const uint4* __restrict__ bits, uint8_t* __restrict__ coarse) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_bytes) return;
	uint32_t out = 0;
	#pragma unroll
	for (uint32_t b = 0; b < 8; ++b) {
		uint32_t any = 0;
		#pragma unroll
		for (uint32_t q = 0; q < 4; ++q) { const uint4 v = bits[((size_t)i * 8 + b) * 4 + q]; any |= v.x | v.y | v.z | v.w; }
		if (any) out |= 1u << b;
	}
	coarse[i] = (uint8_t)out;
}

} // namespace ngp

using namespace ngp;

extern "C" int ngp_mark_untrained_grid(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, uint32_t n_images, const float* d_xforms, float* d_grid) {
	NGP_REQUIRE(cfg && cam && d_xforms && d_grid, "null pointer");
	const uint32_t total = cfg->grid_size * cfg->grid_size * cfg->grid_size * cfg->n_cascades;
	NGP_LAUNCH("mark_untrained_grid_kernel", (cudaStream_t)stream, mark_untrained_grid_kernel<<<grid_for(total, 256), 256, 0, (cudaStream_t)stream>>>(*cfg, *cam, n_images, d_xforms, d_grid));
	return check_launch("mark_untrained_grid");
}

extern "C" int ngp_generate_grid_samples(void* stream, const ngp_sampler_cfg* cfg, uint32_t n, uint64_t rng_state, uint64_t rng_inc, uint32_t step,
                                         const float* d_grid, float thresh, float* d_positions, uint32_t* d_indices) {
	NGP_REQUIRE(cfg && d_grid && d_positions && d_indices, "null pointer");
	if (n == 0) return 0;
	Pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
	NGP_LAUNCH("generate_grid_samples_kernel", (cudaStream_t)stream, generate_grid_samples_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(*cfg, n, 0, n, rng, step, d_grid, thresh, d_positions, d_indices));
	return check_launch("generate_grid_samples");
}

extern "C" int ngp_generate_grid_samples_sorted_part(void* stream, const ngp_sampler_cfg* cfg, uint32_t n, uint64_t rng_state, uint64_t rng_inc, uint32_t step,
                                                     const float* d_grid, float thresh, float* d_positions, uint32_t* d_indices,
                                                     uint32_t part, uint32_t n_parts, uint32_t* slot_begin, uint32_t* slot_end) {
	NGP_REQUIRE(cfg && d_grid && d_positions && d_indices, "null pointer");
	NGP_REQUIRE(n_parts >= 1 && part < n_parts, "bad part");
	uint32_t lo = 0, hi = 0;
	if (slot_begin) *slot_begin = 0;
	if (slot_end) *slot_end = 0;
	if (n == 0) return 0;
	Pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
	const uint32_t cells = cfg->grid_size * cfg->grid_size * cfg->grid_size;
	if ((cells & (cells - 1)) != 0 || cells < 256) {         // the inversion needs a power-of-two cell count: plain order otherwise
		lo = (uint32_t)((uint64_t)n * part / n_parts); hi = (uint32_t)((uint64_t)n * (part + 1) / n_parts);
		if (hi > lo) NGP_LAUNCH("generate_grid_samples_kernel", (cudaStream_t)stream, generate_grid_samples_kernel<<<grid_for(hi - lo, 256), 256, 0, (cudaStream_t)stream>>>(*cfg, n, lo, hi, rng, step, d_grid, thresh, d_positions, d_indices));
	} else {
		uint32_t inv = kCellMul;                             // Newton: the inverse of an odd number modulo 2^32
		for (int k = 0; k < 5; ++k) inv *= 2u - kCellMul * inv;
		const uint32_t mask = cells - 1;
		const uint32_t alpha = inv & mask, beta = (0u - inv * kCellAdd - step * n) & mask;
		const uint32_t chunk = div_up(div_up(cells, 256), kNumSMs * 4) * 256;        // one resident wave of CTAs, each walks a contiguous run of cells
		const uint32_t n_blocks = div_up(cells, chunk);
		const uint32_t b0 = (uint32_t)((uint64_t)n_blocks * part / n_parts), b1 = (uint32_t)((uint64_t)n_blocks * (part + 1) / n_parts);
		// samples whose first candidate cell is below m: the kernel's own closed form, on the host
		const uint32_t K = n / cells, R = n % cells;
		auto below = [&](uint64_t m) -> uint64_t { m = m < cells ? m : cells; return m * K + (R ? m + floor_sum(m, cells, alpha, beta) - floor_sum(m, cells, alpha, (uint64_t)beta + cells - R) : 0); };
		lo = (uint32_t)below((uint64_t)b0 * chunk); hi = (uint32_t)below((uint64_t)b1 * chunk);
		if (b1 > b0) NGP_LAUNCH("generate_grid_samples_sorted_kernel", (cudaStream_t)stream, generate_grid_samples_sorted_kernel<<<b1 - b0, 256, 0, (cudaStream_t)stream>>>(*cfg, n, rng, step, alpha, beta, chunk, b0, d_grid, thresh, d_positions, d_indices));
	}
	if (slot_begin) *slot_begin = lo;
	if (slot_end) *slot_end = hi;
	return check_launch("generate_grid_samples_sorted_part");
}

extern "C" int ngp_generate_grid_samples_sorted(void* stream, const ngp_sampler_cfg* cfg, uint32_t n, uint64_t rng_state, uint64_t rng_inc, uint32_t step,
                                                const float* d_grid, float thresh, float* d_positions, uint32_t* d_indices) {
	return ngp_generate_grid_samples_sorted_part(stream, cfg, n, rng_state, rng_inc, step, d_grid, thresh, d_positions, d_indices, 0, 1, nullptr, nullptr);
}

extern "C" int ngp_splat_grid(void* stream, const ngp_sampler_cfg* cfg, uint32_t n, const uint32_t* d_indices, const void* d_density, uint32_t density_stride,
                              int density_activation, float* d_grid_current) {
	NGP_REQUIRE(cfg && d_indices && d_density && d_grid_current, "null pointer");
	cudaStream_t s = (cudaStream_t)stream;
	if (n) NGP_LAUNCH("splat_grid_kernel", s, splat_grid_kernel<<<grid_for(n, 256), 256, 0, s>>>(n, d_indices, (const __half*)d_density, density_stride, density_activation, cfg->min_step, d_grid_current));
	return check_launch("splat_grid");
}

extern "C" int ngp_ema_grid(void* stream, const ngp_sampler_cfg* cfg, float decay, const float* d_grid_current, float* d_grid) {
	NGP_REQUIRE(cfg && d_grid_current && d_grid, "null pointer");
	const uint32_t total = cfg->grid_size * cfg->grid_size * cfg->grid_size * cfg->n_cascades;
	cudaStream_t s = (cudaStream_t)stream;
	NGP_LAUNCH("ema_grid_kernel", s, ema_grid_kernel<<<grid_for(total, 256), 256, 0, s>>>(total, d_grid_current, decay, d_grid));
	return check_launch("ema_grid");
}

extern "C" int ngp_update_grid(void* stream, const ngp_sampler_cfg* cfg, uint32_t n, const uint32_t* d_indices, const void* d_density, uint32_t density_stride,
                               int density_activation, float decay, float* d_grid_current, float* d_grid) {
	NGP_REQUIRE(cfg && d_indices && d_density && d_grid_current && d_grid, "null pointer");
	const uint32_t total = cfg->grid_size * cfg->grid_size * cfg->grid_size * cfg->n_cascades;
	cudaStream_t s = (cudaStream_t)stream;
	if (cudaMemsetAsync(d_grid_current, 0, (size_t)total * sizeof(float), s) != cudaSuccess) { set_error("ngp_update_grid: memset failed"); return -2; }
	if (n) NGP_LAUNCH("splat_grid_kernel", s, splat_grid_kernel<<<grid_for(n, 256), 256, 0, s>>>(n, d_indices, (const __half*)d_density, density_stride, density_activation, cfg->min_step, d_grid_current));
	NGP_LAUNCH("ema_grid_kernel", s, ema_grid_kernel<<<grid_for(total, 256), 256, 0, s>>>(total, d_grid_current, decay, d_grid));
	return check_launch("update_grid");
}

extern "C" int ngp_update_grid_bitfield(void* stream, const ngp_sampler_cfg* cfg, const float* d_grid, float min_optical_thickness, float* d_mean, uint8_t* d_bitfield) {
	NGP_REQUIRE(cfg && d_grid && d_mean && d_bitfield, "null pointer");
	const uint32_t G = cfg->grid_size, cells = G * G * G;
	NGP_REQUIRE(cells % 64 == 0, "grid_size^3 must be a multiple of 64");
	cudaStream_t s = (cudaStream_t)stream;
	NGP_LAUNCH("grid_mean_partial_kernel", s, grid_mean_partial_kernel<<<kMeanParts, 256, 0, s>>>(d_grid, cells, d_mean + 1));
	NGP_LAUNCH("grid_mean_final_kernel", s, grid_mean_final_kernel<<<1, kMeanParts, 0, s>>>(d_mean + 1, d_mean));
	const uint32_t n_bytes = cells * cfg->n_cascades / 8;
	NGP_LAUNCH("grid_to_bitfield_kernel", s, grid_to_bitfield_kernel<<<grid_for(n_bytes, 256), 256, 0, s>>>(n_bytes, d_grid, d_bitfield, d_mean, min_optical_thickness));
	for (uint32_t level = 1; level < cfg->n_cascades; ++level)
		NGP_LAUNCH("bitfield_max_pool_kernel", s, bitfield_max_pool_kernel<<<grid_for(cells / 64, 256), 256, 0, s>>>(cells / 64, G, d_bitfield + (size_t)cells * (level - 1) / 8, d_bitfield + (size_t)cells * level / 8));
	return check_launch("update_grid_bitfield");
}

extern "C" size_t ngp_coarse_bitfield_bytes(const ngp_sampler_cfg* cfg) {
	if (!cfg) return 0;
	return (size_t)cfg->grid_size * cfg->grid_size * cfg->grid_size / 4096 * cfg->n_cascades;
}

extern "C" int ngp_build_coarse_bitfield(void* stream, const ngp_sampler_cfg* cfg, const uint8_t* d_bitfield, uint8_t* d_coarse) {
	NGP_REQUIRE(cfg && d_bitfield && d_coarse, "null pointer");
	const uint32_t G = cfg->grid_size;
	NGP_REQUIRE((G & (G - 1)) == 0 && G >= 32, "the coarse bitfield needs grid_size to be a power of two >= 32");
	NGP_REQUIRE(((uintptr_t)d_bitfield & 15) == 0, "the cell bitfield must be 16-byte aligned");
	const uint32_t n_bytes = (uint32_t)ngp_coarse_bitfield_bytes(cfg);
	cudaStream_t s = (cudaStream_t)stream;
	NGP_LAUNCH("coarse_bitfield_kernel", s, coarse_bitfield_kernel<<<div_up(n_bytes, 128), 128, 0, s>>>(n_bytes, (const uint4*)d_bitfield, d_coarse));
	return check_launch("build_coarse_bitfield");
}
