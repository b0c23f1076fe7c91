// TEST INFRASTRUCTURE ONLY (see oracle_common.h).  Included by sampler.cpp.
// Occupancy-grid construction: occupancy_sampler.cu:18-205, 660-760.

extern "C" {

// grid_out[C*G^3]: 0 where the Morton cell is visible from any camera, else -1.   (:18-61)
void orc_mark_untrained_grid(uint32_t n_cascades, uint32_t G, uint32_t n_images, int W, int H, float fx, float fy,
                             const float* xforms, float* grid_out) {
	const uint32_t cells = G * G * G;
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)cells * n_cascades; ++i) {
		uint32_t level = (uint32_t)(i / cells), pos_idx = (uint32_t)(i % cells);
		uint32_t x = morton3D_invert(pos_idx >> 0), y = morton3D_invert(pos_idx >> 1), z = morton3D_invert(pos_idx >> 2);
		float s = scalbnf(1.0f, (int)level);
		float px = (((float)x + 0.5f) / (float)G - 0.5f) * s + 0.5f;
		float py = (((float)y + 0.5f) / (float)G - 0.5f) * s + 0.5f;
		float pz = (((float)z + 0.5f) / (float)G - 0.5f) * s + 0.5f;
		float half_resx = W * 0.5f, half_resy = H * 0.5f;
		float voxel_radius = 0.5f * 1.73205080757f * s / (float)G;
		int count = 0;
		for (uint32_t j = 0; j < n_images; ++j) {
			const float* xf = xforms + (size_t)j * 12;
			float lx = px - xf[9], ly = py - xf[10], lz = pz - xf[11];
			// Eigen dot(): p0 + (p1 + p2), both sums contracted by nvcc
			float cx = fmaf(lx, xf[0], fmaf(ly, xf[1], lz * xf[2]));
			float cy = fmaf(lx, xf[3], fmaf(ly, xf[4], lz * xf[5]));
			float cz = fmaf(lx, xf[6], fmaf(ly, xf[7], lz * xf[8]));
			if (cz > 0.f) {
				if (fabsf(cx) - voxel_radius < cz / fx * half_resx && fabsf(cy) - voxel_radius < cz / fy * half_resy) { count++; break; }
			}
		}
		grid_out[i] = count > 0 ? 0.f : -1.f;
	}
}

// (:63-105) positions out [n][3] (AABB-relative), indices out [n]
void orc_generate_grid_samples(uint32_t n, uint64_t rng_state, uint64_t rng_inc, uint32_t step, float aabb_min, float aabb_max,
                               const float* grid_in, uint32_t G, uint32_t n_cascades, float thresh, float* out_pos, uint32_t* out_idx) {
	const uint32_t cells = G * G * G;
	#pragma omp parallel for schedule(static)
	for (int64_t ii = 0; ii < (int64_t)n; ++ii) {
		uint32_t i = (uint32_t)ii;
		Pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
		rng.advance((int64_t)((uint64_t)i * 4));
		uint32_t level = (uint32_t)(rng.next_float() * n_cascades) % n_cascades;
		uint32_t idx = 0;
		for (uint32_t j = 0; j < 10; ++j) {
			idx = ((i + step * n) * 56924617u + j * 19349663u + 96925573u) % cells;
			idx += level * cells;
			if (grid_in[idx] > thresh) break;
		}
		uint32_t pos_idx = idx % cells;
		uint32_t x = morton3D_invert(pos_idx >> 0), y = morton3D_invert(pos_idx >> 1), z = morton3D_invert(pos_idx >> 2);
		float rx = rng.next_float(), ry = rng.next_float(), rz = rng.next_float();
		float s = scalbnf(1.0f, (int)level);
		float px = (((float)x + rx) / (float)G - 0.5f) * s + 0.5f;
		float py = (((float)y + ry) / (float)G - 0.5f) * s + 0.5f;
		float pz = (((float)z + rz) / (float)G - 0.5f) * s + 0.5f;
		float diag = aabb_max - aabb_min;
		out_pos[(size_t)i * 3 + 0] = (px - aabb_min) / diag;
		out_pos[(size_t)i * 3 + 1] = (py - aabb_min) / diag;
		out_pos[(size_t)i * 3 + 2] = (pz - aabb_min) / diag;
		out_idx[i] = idx;
	}
}

// splat-max + decay-max (:107-158).  density_out: [n][stride] fp16, element 0 = raw sigma.
// expf stands in for the device's __expf (tolerance, not bit-exact).
void orc_update_grid(uint32_t n, const uint32_t* indices, const uint16_t* density_out, uint32_t stride, float min_step,
                     float decay, uint32_t n_total, float* grid_current, float* grid) {
	for (uint32_t i = 0; i < n_total; ++i) grid_current[i] = 0.f;
	for (uint32_t i = 0; i < n; ++i) {
		float thickness = expf(h16from(density_out[(size_t)i * stride])) * min_step;
		float& g = grid_current[indices[i]];
		uint32_t a, b; std::memcpy(&a, &g, 4); std::memcpy(&b, &thickness, 4);
		if (b > a) g = thickness;   // uint-ordered atomicMax of the reference
	}
	for (uint32_t i = 0; i < n_total; ++i) {
		float prev = grid[i];
		grid[i] = prev < 0.f ? prev : fmaxf(prev * decay, grid_current[i]);
	}
}

// mean over cascade 0 of max(v,0)/cells  (:737-744); the GPU uses a tree, so this is compared with tolerance
float orc_grid_mean(const float* grid, uint32_t cells) {
	double s = 0.0;
	for (uint32_t i = 0; i < cells; ++i) s += (double)(fmaxf(grid[i], 0.f) / (float)cells);
	return (float)s;
}

// bits (:160-180) then max-pool into coarser cascades (:182-205, 753-759)
void orc_grid_to_bitfield(const float* grid, uint32_t G, uint32_t n_cascades, float mean, float min_optical_thickness, uint8_t* bits) {
	const uint32_t cells = G * G * G;
	const float thresh = std::min(min_optical_thickness, mean);
	for (uint32_t i = 0; i < cells * n_cascades / 8; ++i) {
		uint8_t b = 0;
		for (int j = 0; j < 8; ++j) b |= grid[(size_t)i * 8 + j] > thresh ? (uint8_t)(1u << j) : 0;
		bits[i] = b;
	}
	for (uint32_t level = 1; level < n_cascades; ++level) {
		const uint8_t* prev = bits + (size_t)cells * (level - 1) / 8;
		uint8_t* next = bits + (size_t)cells * level / 8;
		for (uint32_t i = 0; i < cells / 64; ++i) {
			uint8_t b = 0;
			for (int j = 0; j < 8; ++j) b |= prev[(size_t)i * 8 + j] > 0 ? (uint8_t)(1u << j) : 0;
			uint32_t x = morton3D_invert(i >> 0) + G / 8, y = morton3D_invert(i >> 1) + G / 8, z = morton3D_invert(i >> 2) + G / 8;
			next[morton3D(x, y, z)] |= b;
		}
	}
}

} // extern "C"
