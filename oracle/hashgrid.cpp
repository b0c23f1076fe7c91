// TEST INFRASTRUCTURE ONLY (see oracle_common.h).  Pinned to golden vectors from the reference's CUDA build (tests/golden/).
// CPU restatement of the multiresolution hash-grid encoding.
//   offsets / resolutions : tiny-cuda-nn/encodings/grid.h:475-537
//   forward               : grid.h:122-215 (kernel_grid), pos_fract common_device.h:396-413
//   index                 : grid.h:66-100  (fast_hash, grid_index)
//   backward scatter      : grid.h:259-368 (kernel_grid_backward)
#include "oracle_common.h"
#include <omp.h>

using namespace orc;

extern "C" {

// per_level_scale when the config leaves it out (internal/network/ngp_network.h:623-632), float math
float orc_auto_per_level_scale(float desired_resolution, uint32_t base_resolution, uint32_t n_levels) {
	return std::exp(std::log(desired_resolution / (float)base_resolution) / (n_levels - 1));
}

// Level table. scale[l] = exp2f(l*log2(per_level_scale))*base - 1 (host float math as
// in the ctor, grid.h:498); resolution = ceil(scale)+1; size = min(next_multiple(res^3,8), 2^log2T)
// offsets in entries (n_levels+1 values).  Returns total entries.
uint32_t orc_grid_levels(uint32_t n_levels, uint32_t log2_hashmap_size, uint32_t base_resolution,
                         float per_level_scale, uint32_t* offsets, float* scales, uint32_t* resolutions) {
	uint32_t offset = 0;
	for (uint32_t i = 0; i < n_levels; ++i) {
		const float scale = exp2f(i * std::log2(per_level_scale)) * base_resolution - 1.0f;
		const uint32_t res = (uint32_t)(ceilf(scale)) + 1;
		uint32_t max_params = 0xFFFFFFFFu / 2;
		uint32_t params = std::pow((float)res, 3) > (float)max_params ? max_params : res * res * res;
		params = next_multiple(params, 8u);
		params = std::min(params, 1u << log2_hashmap_size);
		offsets[i] = offset;
		if (scales) scales[i] = scale;
		if (resolutions) resolutions[i] = res;
		offset += params;
	}
	offsets[n_levels] = offset;
	return offset;
}

static inline uint32_t grid_index(uint32_t size, uint32_t res, uint32_t gx, uint32_t gy, uint32_t gz) {
	uint32_t stride = 1, index = 0;
	const uint32_t g[3] = {gx, gy, gz};
	for (uint32_t d = 0; d < 3 && stride <= size; ++d) { index += g[d] * stride; stride *= res; }
	if (size < stride) index = (gx * 1u) ^ (gy * 19349663u) ^ (gz * 83492791u);
	return index % size;
}

// entry index of one corner; exposed for known-answer tests
uint32_t orc_grid_index(uint32_t size, uint32_t res, uint32_t gx, uint32_t gy, uint32_t gz) {
	return grid_index(size, res, gx, gy, gz);
}

// xyz: n points, stride_floats apart.  table: fp16 bit patterns, 2 features / entry.
// out: [n][2*n_levels] fp16 bits.  indices (optional): [n][n_levels][8] u32 entry index within level.
void orc_hash_encode_fwd(const float* xyz, uint32_t stride_floats, uint32_t n,
                         const uint16_t* table, uint32_t n_levels, const uint32_t* offsets,
                         const float* scales, const uint32_t* resolutions,
                         uint16_t* out, uint32_t* indices) {
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)n; ++i) {
		const float* p = xyz + (size_t)i * stride_floats;
		for (uint32_t l = 0; l < n_levels; ++l) {
			const uint32_t size = offsets[l + 1] - offsets[l];
			const uint16_t* grid = table + (size_t)offsets[l] * 2;
			const float scale = scales[l];
			const uint32_t res = resolutions[l];
			float w[3]; uint32_t g[3];
			for (int d = 0; d < 3; ++d) {
				float pos = fmaf(p[d], scale, 0.5f);
				int tmp = (int)floorf(pos);
				g[d] = (uint32_t)tmp;
				w[d] = pos - (float)tmp;
			}
			float acc0 = 0.f, acc1 = 0.f;
			for (uint32_t c = 0; c < 8; ++c) {
				float weight = 1.f; uint32_t gl[3];
				for (int d = 0; d < 3; ++d) {
					if ((c & (1u << d)) == 0) { weight *= 1 - w[d]; gl[d] = g[d]; }
					else { weight *= w[d]; gl[d] = g[d] + 1; }
				}
				uint32_t e = grid_index(size, res, gl[0], gl[1], gl[2]);
				if (indices) indices[((size_t)i * n_levels + l) * 8 + c] = e;
				float v0 = h16from(grid[(size_t)e * 2 + 0]);
				float v1 = h16from(grid[(size_t)e * 2 + 1]);
				acc0 = h16add(h16r(weight * v0), acc0);
				acc1 = h16add(h16r(weight * v1), acc1);
			}
			out[(size_t)i * n_levels * 2 + l * 2 + 0] = h16bits(acc0);
			out[(size_t)i * n_levels * 2 + l * 2 + 1] = h16bits(acc1);
		}
	}
}

// Deterministic backward: the reference sums fp16 terms with fp16 atomics in arbitrary order;
// the oracle sums the same fp16-rounded terms in fp32 (grad_f32, zero-initialised by the caller,
// 2 floats / entry).  dy: [n][2*n_levels] fp16 bits.
void orc_hash_encode_bwd(const float* xyz, uint32_t stride_floats, uint32_t n,
                         const uint16_t* dy, uint32_t n_levels, const uint32_t* offsets,
                         const float* scales, const uint32_t* resolutions, float* grad_f32) {
	// parallel over levels: levels own disjoint table slices, so no atomics are needed
	#pragma omp parallel for schedule(dynamic, 1)
	for (int l = 0; l < (int)n_levels; ++l) {
		const uint32_t size = offsets[l + 1] - offsets[l];
		float* grid = grad_f32 + (size_t)offsets[l] * 2;
		const float scale = scales[l];
		const uint32_t res = resolutions[l];
		for (uint32_t i = 0; i < n; ++i) {
			const float* p = xyz + (size_t)i * stride_floats;
			float w[3]; uint32_t g[3];
			for (int d = 0; d < 3; ++d) {
				float pos = fmaf(p[d], scale, 0.5f);
				int tmp = (int)floorf(pos);
				g[d] = (uint32_t)tmp;
				w[d] = pos - (float)tmp;
			}
			const float g0 = h16from(dy[(size_t)i * n_levels * 2 + l * 2 + 0]);
			const float g1 = h16from(dy[(size_t)i * n_levels * 2 + l * 2 + 1]);
			for (uint32_t c = 0; c < 8; ++c) {
				float weight = 1.f; uint32_t gl[3];
				for (int d = 0; d < 3; ++d) {
					if ((c & (1u << d)) == 0) { weight *= 1 - w[d]; gl[d] = g[d]; }
					else { weight *= w[d]; gl[d] = g[d] + 1; }
				}
				uint32_t e = grid_index(size, res, gl[0], gl[1], gl[2]);
				grid[(size_t)e * 2 + 0] += h16r(g0 * weight);
				grid[(size_t)e * 2 + 1] += h16r(g1 * weight);
			}
		}
	}
}

} // extern "C"
