// Multiresolution hash-grid encoding, forward gather and backward scatter, for sm_100a.
// Replaces tiny-cuda-nn/encodings/grid.h: extract_position (:108), kernel_grid (:122), transpose_encoded_position (:371),
// transpose_gradients (:386), kernel_grid_backward (:259) -- one kernel per direction, no SoA/AoS round trips.
//
// Layout: a warp owns 32 consecutive samples and walks the levels; consecutive samples along a ray share cells on the
// coarse levels, so a warp's gathers hit few cache lines and its scatter can be reduced in-warp before the atomic.
// The whole table (23 MB at T=2^19) stays resident in B200's 126 MB L2, so levels need not be swept one at a time.
// Roofline: L2 bandwidth (588 algorithmic bytes / sample).
#include "common.cuh"

namespace ngp {

struct LevelInfo { uint32_t offset, size, res; float scale; };

__device__ __forceinline__ uint32_t grid_index(uint32_t size, uint32_t res, bool hashed, uint32_t gx, uint32_t gy, uint32_t gz) {
	// grid.h:84-100 -- dense index while the stride fits the table, spatial hash otherwise
	uint32_t index;
	if (hashed) {
		index = gx ^ (gy * 19349663u) ^ (gz * 83492791u);
	} else {
		uint32_t stride = 1; index = 0;
		index += gx * stride; stride *= res;
		if (stride <= size) { index += gy * stride; stride *= res; }
		if (stride <= size) { index += gz * stride; }
	}
	return (size & (size - 1)) == 0 ? (index & (size - 1)) : (index % size);
}

__device__ __forceinline__ bool level_is_hashed(uint32_t size, uint32_t res) {
	// grid.h:88-95: the stride is a uint32_t and is multiplied once more after the last dimension; "size < stride" then decides.
	// The wrap-around is part of the contract: with T = 2^22 the finest level has res^2 == size, the third product is 2^33 = 0
	// (mod 2^32), and the reference indexes that level DENSELY as (gx + gy * res + gz * res^2) mod size -- found by the T = 2^22
	// parity test, a 64-bit stride here called the level hashed.
	uint32_t s = 1;
	for (int d = 0; d < 3 && s <= size; ++d) s *= res;
	return size < s;
}

__device__ __forceinline__ void pos_fract(float x, float scale, uint32_t& g, float& w) {
	float p = __fmaf_rn(x, scale, 0.5f);
	float f = floorf(p);
	g = (uint32_t)(int)f;
	w = p - f;
}

// x-adjacent corners share one aligned 16-byte load whenever both fall into one group of four entries (3 cases out of 4).  The kernel
// is bound by the tag stage of L1TEX -- one look-up per lane and instruction once every lane hits its own line -- so what counts
// is the number of gather instructions per level (8 with 4-byte loads, 6 with 8-byte pairs, 5 here), not the bytes they return.
template <bool WRITE_INDICES>
__global__ void __launch_bounds__(256) hash_encode_fwd_kernel(
	const float* __restrict__ xyz, uint32_t stride_floats, uint32_t n, const uint32_t* __restrict__ n_valid,
	const __half2* __restrict__ table, const __grid_constant__ ngp_grid_desc desc,
	__half* __restrict__ out, uint32_t out_stride, uint32_t* __restrict__ indices)
{
	chain_enter();
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (n_valid && blockIdx.x * blockDim.x >= *n_valid) return;
	if (i >= n) return;
	const float x = xyz[(size_t)i * stride_floats + 0], y = xyz[(size_t)i * stride_floats + 1], z = xyz[(size_t)i * stride_floats + 2];
	__half* orow = out + (size_t)i * out_stride;

	#pragma unroll 2
	for (uint32_t l = 0; l < desc.n_levels; ++l) {
		const uint32_t off = desc.offsets[l], size = desc.offsets[l + 1] - off, res = desc.resolutions[l];
		const float scale = desc.scales[l];
		const bool hashed = level_is_hashed(size, res);
		const __half2* grid = table + off;
		uint32_t gx, gy, gz; float wx, wy, wz;
		pos_fract(x, scale, gx, wx); pos_fract(y, scale, gy, wy); pos_fract(z, scale, gz, wz);
		uint32_t e[8];
		#pragma unroll
		for (int c = 0; c < 8; ++c)
			e[c] = grid_index(size, res, hashed, gx + (c & 1), gy + ((c >> 1) & 1), gz + ((c >> 2) & 1));
		// The two x-neighbours of a corner pair are adjacent table entries whenever the lower one has an even index (hashed:
		// gx ^ h and (gx+1) ^ h differ in bit 0 only when gx is even; dense: e and e+1).  One aligned 8-byte load then brings
		// both, and only the other half of the lanes issue a second gather: a quarter fewer L1 wavefronts per level.
		__half2 v[8];
		{
			uint4 q[4];
			#pragma unroll
			for (int c = 0; c < 8; c += 2) q[c >> 1] = __ldg(reinterpret_cast<const uint4*>(grid) + (e[c] >> 2));
			#pragma unroll
			for (int c = 0; c < 8; c += 2) {
				const uint4 t = q[c >> 1];
				const uint32_t k0 = e[c] & 3u, k1 = e[c + 1] & 3u;
				const uint32_t a = (k0 & 2u) ? ((k0 & 1u) ? t.w : t.z) : ((k0 & 1u) ? t.y : t.x);
				v[c] = *reinterpret_cast<const __half2*>(&a);
				if ((e[c + 1] >> 2) == (e[c] >> 2)) {
					const uint32_t b = (k1 & 2u) ? ((k1 & 1u) ? t.w : t.z) : ((k1 & 1u) ? t.y : t.x);
					v[c + 1] = *reinterpret_cast<const __half2*>(&b);
				} else v[c + 1] = __ldg(grid + e[c + 1]);
			}
		}
		__half2 acc = __floats2half2_rn(0.f, 0.f);
		#pragma unroll
		for (int c = 0; c < 8; ++c) {
			float w = 1.f;
			w *= (c & 1) ? wx : 1 - wx;
			w *= (c & 2) ? wy : 1 - wy;
			w *= (c & 4) ? wz : 1 - wz;
			const float2 f = __half22float2(v[c]);
			// each term rounded to fp16, sum carried in fp16 (grid.h:205-212)
			acc = __hadd2(__floats2half2_rn(w * f.x, w * f.y), acc);
			if (WRITE_INDICES) indices[((size_t)i * desc.n_levels + l) * 8 + c] = e[c];
		}
		*reinterpret_cast<__half2*>(orow + 2 * l) = acc;
	}
}

// all 8 corner entries of one cell; same values as grid_index() corner by corner.  Hashed: the y and z products are formed once
// and stepped by their primes.  Dense: one base index stepped by the strides, wrapped by a compare (the index of a point inside
// the unit cube is below 2 * size), with the general modulo kept for anything else.
__device__ __forceinline__ void cell_entries(uint32_t size, uint32_t res, bool hashed, uint32_t gx, uint32_t gy, uint32_t gz, uint32_t (&e)[8]) {
	if (hashed) {
		const uint32_t hy0 = gy * 19349663u, hy1 = hy0 + 19349663u, hz0 = gz * 83492791u, hz1 = hz0 + 83492791u;
		const uint32_t h[4] = {hy0 ^ hz0, hy1 ^ hz0, hy0 ^ hz1, hy1 ^ hz1};
		const bool pow2 = (size & (size - 1)) == 0;
		#pragma unroll
		for (int c = 0; c < 8; ++c) {
			const uint32_t idx = (gx + (c & 1)) ^ h[c >> 1];
			e[c] = pow2 ? (idx & (size - 1)) : (idx % size);
		}
	} else {
		const uint32_t sy = res, sz = res * res;
		const uint32_t base = gx + gy * sy + gz * sz;
		#pragma unroll
		for (int c = 0; c < 8; ++c) {
			uint32_t idx = base + (c & 1) + ((c & 2) ? sy : 0u) + ((c & 4) ? sz : 0u);
			if (idx >= size) { idx -= size; if (idx >= size) idx = (size & (size - 1)) == 0 ? (idx & (size - 1)) : (idx % size); }
			e[c] = idx;
		}
	}
}

__device__ __forceinline__ void red_h2(__half2* p, __half2 v) {
	asm volatile("red.global.add.noftz.f16x2 [%0], %1;" :: "l"(p), "r"(*reinterpret_cast<const uint32_t*>(&v)) : "memory");
}
__device__ __forceinline__ void red_h2x2(__half2* p, __half2 a, __half2 b) {
	asm volatile("red.global.add.noftz.v2.f16x2 [%0], {%1, %2};" :: "l"(p), "r"(*reinterpret_cast<const uint32_t*>(&a)), "r"(*reinterpret_cast<const uint32_t*>(&b)) : "memory");
}

// one 16-byte reduction onto four adjacent half2 entries (p 16-byte aligned); unused slots carry -0, the identity of the sum
__device__ __forceinline__ void red_h2x4(__half2* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
	asm volatile("red.global.add.noftz.v4.f16x2 [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// Backward.  Same mapping (thread = sample, warp = 32 consecutive samples of mostly one ray) and the same
// arithmetic per term (weight product order, fp16 rounding of every term as in grid.h:303-306), but the merge of equal targets
// is keyed by the CELL: neighbouring lanes inside one cell share all 8 corners, so the run structure is found once per level
// (3 shuffles + 1 ballot) instead of once per corner, the 16 partial sums of a run are reduced in fp32 with only as many
// shuffle rounds as the longest run of the warp needs, and warps without any run (the fine levels) skip the merge altogether.
// x-adjacent corner pairs whose entries are adjacent in memory leave as one 8- or 16-byte reduction.  (A first-generation kernel
// keyed the merge per corner: ~800 instructions per warp and level, issue-bound at 72 %, 144.8 us against 78 us -- DESIGN.md 3.1.)
template <bool QUADS>
__global__ void __launch_bounds__(256) hash_encode_bwd_cell_kernel(
	const float* __restrict__ xyz, uint32_t stride_floats, uint32_t n,
	const __half* __restrict__ dy, uint32_t dy_stride, const __grid_constant__ ngp_grid_desc desc, __half2* __restrict__ grad,
	uint32_t level_begin, uint32_t level_end)
{
	chain_enter();
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t lane = threadIdx.x & 31;
	const bool valid = i < n;
	const uint32_t ii = valid ? i : n - 1;
	const float x = xyz[(size_t)ii * stride_floats + 0], y = xyz[(size_t)ii * stride_floats + 1], z = xyz[(size_t)ii * stride_floats + 2];
	const __half* drow = dy + (size_t)ii * dy_stride;

	for (uint32_t l = level_begin; l < level_end; ++l) {
		const uint32_t off = desc.offsets[l], size = desc.offsets[l + 1] - off, res = desc.resolutions[l];
		const float scale = desc.scales[l];
		const bool hashed = level_is_hashed(size, res);
		__half2* grid = grad + off;
		uint32_t gx, gy, gz; float wx, wy, wz;
		pos_fract(x, scale, gx, wx); pos_fract(y, scale, gy, wy); pos_fract(z, scale, gz, wz);
		const float2 g = __half22float2(*reinterpret_cast<const __half2*>(drow + 2 * l));

		// runs of neighbouring lanes inside one cell (invalid lanes sit at the tail of the last warp and are runs of their own)
		const uint32_t pgx = __shfl_up_sync(0xffffffffu, gx, 1), pgy = __shfl_up_sync(0xffffffffu, gy, 1), pgz = __shfl_up_sync(0xffffffffu, gz, 1);
		const bool joins = valid && lane != 0 && pgx == gx && pgy == gy && pgz == gz;
		const uint32_t heads = ~__ballot_sync(0xffffffffu, joins);

		uint32_t e[8];
		cell_entries(size, res, hashed, gx, gy, gz, e);
		__half2 t[8];
		#pragma unroll
		for (int c = 0; c < 8; ++c) {
			// same product order as the forward pass: (wx' * wy') * wz'
			const float w = (((c & 1) ? wx : 1 - wx) * ((c & 2) ? wy : 1 - wy)) * ((c & 4) ? wz : 1 - wz);
			t[c] = __floats2half2_rn(g.x * w, g.y * w);
		}
		bool head = true;
		if (heads != 0xffffffffu) {
			const uint32_t above = heads & ~((2u << lane) - 1u);              // heads strictly above this lane
			const uint32_t end = above ? (uint32_t)__ffs(above) - 2u : 31u;     // last lane of this lane's run
			const uint32_t reach = __reduce_max_sync(0xffffffffu, end - lane);  // longest run of the warp, minus one
			float v[16];
			#pragma unroll
			for (int c = 0; c < 8; ++c) { const float2 f = __half22float2(t[c]); v[2 * c] = f.x; v[2 * c + 1] = f.y; }
			for (uint32_t o = 1; o <= reach; o <<= 1) {
				const bool take = lane + o <= end;
				#pragma unroll
				for (int k = 0; k < 16; ++k) { const float a = __shfl_down_sync(0xffffffffu, v[k], o); if (take) v[k] += a; }
			}
			#pragma unroll
			for (int c = 0; c < 8; ++c) t[c] = __floats2half2_rn(v[2 * c], v[2 * c + 1]);
			head = (heads >> lane) & 1u;
		}
		if (valid && head) {
			#pragma unroll
			for (int c = 0; c < 8; c += 2) {
				if (QUADS && (e[c + 1] >> 2) == (e[c] >> 2) && e[c + 1] != e[c]) {
					// both entries inside one aligned group of four: one 16-byte reduction (3 cases out of 4 instead of 2 out of 4)
					const uint32_t k0 = e[c] & 3u, k1 = e[c + 1] & 3u, a = *reinterpret_cast<const uint32_t*>(&t[c]), b = *reinterpret_cast<const uint32_t*>(&t[c + 1]);
					const uint32_t nz = 0x80008000u;
					red_h2x4(grid + (e[c] & ~3u), k0 == 0 ? a : (k1 == 0 ? b : nz), k0 == 1 ? a : (k1 == 1 ? b : nz), k0 == 2 ? a : (k1 == 2 ? b : nz), k0 == 3 ? a : (k1 == 3 ? b : nz));
				} else if (e[c + 1] == (e[c] ^ 1u)) {
					if (e[c] & 1u) red_h2x2(grid + e[c + 1], t[c + 1], t[c]);
					else red_h2x2(grid + e[c], t[c], t[c + 1]);
				} else {
					red_h2(grid + e[c], t[c]);
					red_h2(grid + e[c + 1], t[c + 1]);
				}
			}
		}
	}
}

// dL/dx of the encoding (grid.h:217-255 dy_dx + kernel_grid_backward_input :400-422, fused: dy_dx is never stored).  Per level and
// input dimension the derivative of the trilinear blend is scale * sum over the 4 corner pairs along that dimension of
// (product of the other two weights) * (right - left); dL/dx_j = sum_k dL/dy_k * dy_k/dx_j in feature order k = 2 * level + f.
// The result overwrites the sample's Coordinate: position <- gradient * inv_scale, dt <- dt * inv_scale, direction <- 0.
__global__ void __launch_bounds__(256) hash_input_grad_kernel(float* __restrict__ coords, uint32_t n, const uint32_t* __restrict__ n_valid,
                                                              const __half2* __restrict__ table, const __grid_constant__ ngp_grid_desc desc,
                                                              const __half* __restrict__ denc, float inv_scale) {
	chain_enter();
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= (n_valid ? min(*n_valid, n) : n)) return;
	float* co = coords + (size_t)i * 7;
	const float pos[3] = {co[0], co[1], co[2]};
	const __half2* drow = reinterpret_cast<const __half2*>(denc + (size_t)i * 32);
	float g[3] = {0.f, 0.f, 0.f};
	for (uint32_t l = 0; l < desc.n_levels; ++l) {
		const uint32_t off = desc.offsets[l], size = desc.offsets[l + 1] - off, res = desc.resolutions[l];
		const float scale = desc.scales[l];
		const bool hashed = level_is_hashed(size, res);
		uint32_t gi[3]; float w[3];
		pos_fract(pos[0], scale, gi[0], w[0]); pos_fract(pos[1], scale, gi[1], w[1]); pos_fract(pos[2], scale, gi[2], w[2]);
		uint32_t e[8];
		cell_entries(size, res, hashed, gi[0], gi[1], gi[2], e);
		float2 v[8];
		#pragma unroll
		for (int c = 0; c < 8; ++c) v[c] = __half22float2(__ldg(table + off + e[c]));
		const float2 d = __half22float2(drow[l]);
		#pragma unroll
		for (int gd = 0; gd < 3; ++gd) {
			float gx = 0.f, gy = 0.f;
			#pragma unroll
			for (int idx = 0; idx < 4; ++idx) {
				float weight = scale;
				int corner = 0;
				#pragma unroll
				for (int ngd = 0; ngd < 2; ++ngd) {
					const int dim = ngd >= gd ? ngd + 1 : ngd;
					if ((idx & (1 << ngd)) == 0) weight *= 1 - w[dim];
					else { weight *= w[dim]; corner |= 1 << dim; }
				}
				const float2 left = v[corner], right = v[corner | (1 << gd)];
				gx += weight * (right.x - left.x);
				gy += weight * (right.y - left.y);
			}
			g[gd] += d.x * gx;
			g[gd] += d.y * gy;
		}
	}
	co[0] = g[0] * inv_scale; co[1] = g[1] * inv_scale; co[2] = g[2] * inv_scale;
	co[3] = co[3] * inv_scale;
	co[4] = 0.f; co[5] = 0.f; co[6] = 0.f;
}

int hash_input_gradient(cudaStream_t stream, float* d_coords, uint32_t n, const uint32_t* d_n_valid, const void* d_table, const ngp_grid_desc* desc, const void* d_denc, float inv_scale) {
	if (n == 0) return 0;
	NGP_LAUNCH("hash_input_grad_kernel", stream, launch_chained(hash_input_grad_kernel, div_up(n, 256), 256, 0, stream, d_coords, n, d_n_valid, (const __half2*)d_table, *desc, (const __half*)d_denc, inv_scale));
	return check_launch("hash_input_gradient");
}

// The reference's kernels do not read a scale table: every thread evaluates scale = exp2f(level * log2_per_level_scale) *
// base_resolution - 1 ON THE DEVICE (grid.h:157, :293), i.e. MUFU.EX2 and one contracted FFMA, and derives the resolution from it.
// glibc's exp2f gives a float that differs by an ulp on some levels (2, 5, 6, 8, 10 of base.json), which moves 0.3-5 % of the
// encodings of those levels by one fp16 ulp -- found by running the reference's own GridEncoding beside ours (tests/c1_bench.py).
// This kernel evaluates the same expression once per level; ngp_grid_desc_device_scales stores the results in the descriptor.
__global__ void grid_scales_kernel(uint32_t n_levels, float log2_per_level_scale, uint32_t base_resolution, float* __restrict__ scales) {
	const uint32_t level = threadIdx.x;
	if (level < n_levels) scales[level] = __fmaf_rn(exp2f((float)level * log2_per_level_scale), (float)base_resolution, -1.0f);
}

// zeroes the gradient table as a link of the step's launch chain (a cudaMemsetAsync in the middle of the chain costs two full
// launch boundaries); also leaves the table in L2 for the reductions that follow
__global__ void __launch_bounds__(256) zero_fill_kernel(uint4* __restrict__ p, uint64_t n16, uint8_t* __restrict__ tail, uint32_t n_tail) {
	chain_enter();
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n16; i += (uint64_t)gridDim.x * blockDim.x) p[i] = make_uint4(0u, 0u, 0u, 0u);
	if (blockIdx.x == 0 && threadIdx.x < n_tail) tail[threadIdx.x] = 0;
}

int zero_fill(cudaStream_t stream, void* d_ptr, size_t bytes) {
	if (bytes == 0) return 0;
	if (((uintptr_t)d_ptr & 15u) != 0) {
		if (cudaMemsetAsync(d_ptr, 0, bytes, stream) != cudaSuccess) { set_error("zero_fill: memset failed"); return -2; }
		return 0;
	}
	const uint64_t n16 = bytes / 16;
	NGP_LAUNCH("zero_fill_kernel", stream, launch_chained(zero_fill_kernel, grid_for(n16 ? n16 : 1, 256, 8), 256, 0, stream, (uint4*)d_ptr, n16, (uint8_t*)d_ptr + n16 * 16, (uint32_t)(bytes & 15u)));
	return check_launch("zero_fill");
}

} // namespace ngp

using namespace ngp;

extern "C" float ngp_auto_per_level_scale(float desired_resolution, uint32_t base_resolution, uint32_t n_levels) {
	return n_levels > 1 ? std::exp(std::log(desired_resolution / (float)base_resolution) / (n_levels - 1)) : 1.0f;
}

extern "C" uint32_t ngp_grid_desc_init(ngp_grid_desc* desc, uint32_t n_levels, uint32_t log2_hashmap_size, uint32_t base_resolution, float per_level_scale) {
	if (!desc || n_levels == 0 || n_levels > NGP_MAX_LEVELS) { set_error("ngp_grid_desc_init: n_levels must be in [1,%d]", NGP_MAX_LEVELS); return 0; }
	memset(desc, 0, sizeof(*desc));
	desc->n_levels = n_levels;
	uint32_t offset = 0;
	for (uint32_t i = 0; i < n_levels; ++i) {
		const float scale = exp2f(i * std::log2(per_level_scale)) * base_resolution - 1.0f;
		const uint32_t res = (uint32_t)(ceilf(scale)) + 1;
		const uint32_t max_params = 0xFFFFFFFFu / 2;
		uint32_t params = std::pow((float)res, 3) > (float)max_params ? max_params : res * res * res;
		params = ((params + 7u) / 8u) * 8u;
		params = params < (1u << log2_hashmap_size) ? params : (1u << log2_hashmap_size);
		desc->offsets[i] = offset; desc->scales[i] = scale; desc->resolutions[i] = res;
		offset += params;
	}
	desc->offsets[n_levels] = offset;
	return offset;
}

extern "C" int ngp_grid_desc_device_scales(ngp_grid_desc* desc, uint32_t base_resolution, float per_level_scale, void* stream, float* d_scratch) {
	NGP_REQUIRE(desc && desc->n_levels >= 1 && desc->n_levels <= NGP_MAX_LEVELS, "bad descriptor");
	NGP_REQUIRE(d_scratch, "d_scratch (NGP_MAX_LEVELS floats of device memory) is required");
	cudaStream_t s = (cudaStream_t)stream;
	grid_scales_kernel<<<1, NGP_MAX_LEVELS, 0, s>>>(desc->n_levels, std::log2(per_level_scale), base_resolution, d_scratch);
	float h[NGP_MAX_LEVELS];
	cudaError_t e = cudaMemcpyAsync(h, d_scratch, sizeof(float) * NGP_MAX_LEVELS, cudaMemcpyDeviceToHost, s);
	if (e == cudaSuccess) e = cudaStreamSynchronize(s);          // set-up call: the one entry point that waits for its own result
	if (e != cudaSuccess) { set_error("ngp_grid_desc_device_scales: %s", cudaGetErrorString(e)); return -2; }
	for (uint32_t l = 0; l < desc->n_levels; ++l) {
		desc->scales[l] = h[l];
		desc->resolutions[l] = (uint32_t)ceilf(h[l]) + 1;      // grid.h:158, from the device value; the offsets keep the host value (grid.h:497-510)
	}
	return 0;
}

extern "C" int ngp_hash_encode_fwd(void* stream, const float* d_xyz, uint32_t stride_floats, uint32_t n, const void* d_table,
                                   const ngp_grid_desc* desc, void* d_out, uint32_t out_stride, uint32_t* d_indices) {
	NGP_REQUIRE(d_xyz && d_table && desc && d_out, "null pointer");
	NGP_REQUIRE(((uintptr_t)d_table & 15u) == 0, "d_table must be 16-byte aligned");
	NGP_REQUIRE(out_stride >= 2 * desc->n_levels && out_stride % 2 == 0, "out_stride too small or odd");
	if (n == 0) return 0;
	const uint32_t grid = div_up(n, 256);
	if (d_indices) NGP_LAUNCH("hash_encode_fwd_kernel", (cudaStream_t)stream, launch_chained(hash_encode_fwd_kernel<true>, grid, 256, 0, (cudaStream_t)stream, d_xyz, stride_floats, n, nullptr, (const __half2*)d_table, *desc, (__half*)d_out, out_stride, d_indices));
	else NGP_LAUNCH("hash_encode_fwd_kernel", (cudaStream_t)stream, launch_chained(hash_encode_fwd_kernel<false>, grid, 256, 0, (cudaStream_t)stream, d_xyz, stride_floats, n, nullptr, (const __half2*)d_table, *desc, (__half*)d_out, out_stride, nullptr));
	return check_launch("hash_encode_fwd");
}

namespace ngp {
int hash_encode_fwd_counted(cudaStream_t stream, const float* d_xyz, uint32_t stride_floats, uint32_t n, const uint32_t* d_n_valid,
                            const void* d_table, const ngp_grid_desc* desc, void* d_out, uint32_t out_stride) {
	if (n == 0) return 0;
	// (an explicit L2 prefetch of the 24 MB table before this launch was measured: no change, the kernel is L1TEX-bound, not
	// bound by its first-touch DRAM misses)
	if (((uintptr_t)d_table & 15u) != 0) { set_error("hash_encode_fwd: d_table must be 16-byte aligned"); return -1; }
	NGP_LAUNCH("hash_encode_fwd_kernel", stream, launch_chained(hash_encode_fwd_kernel<false>, div_up(n, 256), 256, 0, stream, d_xyz, stride_floats, n, d_n_valid, (const __half2*)d_table, *desc, (__half*)d_out, out_stride, nullptr));
	return check_launch("hash_encode_fwd");
}
}

namespace ngp {
// the scatter restricted to the levels [level_begin, level_end), accumulating into d_grad (never zeroes it)
int hash_encode_bwd_levels(cudaStream_t stream, const float* d_xyz, uint32_t stride_floats, uint32_t n, const void* d_dy, uint32_t dy_stride,
                           const ngp_grid_desc* desc, void* d_grad, uint32_t level_begin, uint32_t level_end) {
	if (((uintptr_t)d_grad & 7u) != 0) { set_error("hash_encode_bwd: d_grad must be 8-byte aligned"); return -1; }
	if (!(level_begin <= level_end && level_end <= desc->n_levels)) { set_error("hash_encode_bwd: bad level range"); return -1; }
	if (n == 0 || level_begin == level_end) return 0;
	// 16-byte reductions (REDG.E.ADD.F16x8) when the arena is 16-byte aligned, 8-byte pairs otherwise
	if (((uintptr_t)d_grad & 15u) == 0) NGP_LAUNCH("hash_encode_bwd_kernel", stream, launch_chained(hash_encode_bwd_cell_kernel<true>, div_up(n, 256), 256, 0, stream, d_xyz, stride_floats, n, (const __half*)d_dy, dy_stride, *desc, (__half2*)d_grad, level_begin, level_end));
	else NGP_LAUNCH("hash_encode_bwd_kernel", stream, launch_chained(hash_encode_bwd_cell_kernel<false>, div_up(n, 256), 256, 0, stream, d_xyz, stride_floats, n, (const __half*)d_dy, dy_stride, *desc, (__half2*)d_grad, level_begin, level_end));
	return check_launch("hash_encode_bwd");
}
}

extern "C" int ngp_hash_encode_bwd(void* stream, const float* d_xyz, uint32_t stride_floats, uint32_t n, const void* d_dy, uint32_t dy_stride,
                                   const ngp_grid_desc* desc, void* d_grad, int zero_first) {
	NGP_REQUIRE(d_xyz && d_dy && desc && d_grad, "null pointer");
	if (zero_first) {
		const int rc = zero_fill((cudaStream_t)stream, d_grad, (size_t)desc->offsets[desc->n_levels] * 2 * sizeof(__half));
		if (rc) return rc;
	}
	return hash_encode_bwd_levels((cudaStream_t)stream, d_xyz, stride_floats, n, d_dy, dy_stride, desc, d_grad, 0, desc->n_levels);
}
