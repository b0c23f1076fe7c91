// TEST INFRASTRUCTURE ONLY (see oracle_common.h).  Included by render.cpp.
// Inference compositing (ray_marcher.cu:215-295), shade (testbed.cu:448-477),
// accumulate / tonemap (render_buffer.cu:51-186).

extern "C" {

// payloads [n_rays] (40 B), rgba [n_rays][4] f32 in/out.  cfg_f: min_step, T_thresh, depth_scale, aabb_min, aabb_max,
// cam_fwd[3] (camera_matrix col 2).  cfg_u: n_cascades, rgb_act, density_act, render_mode
void orc_composite_infer(uint32_t n_rays, uint32_t n_steps, uint32_t n_steps_offset, void* payloads_, float* rgba,
                         const float* coords, const uint16_t* net_out, uint32_t padded_width,
                         const float* cfg_f, const uint32_t* cfg_u) {
	RayPayload* payloads = (RayPayload*)payloads_;
	const float min_step = cfg_f[0], T_thresh = cfg_f[1], depth_scale = cfg_f[2], amin = cfg_f[3], amax = cfg_f[4];
	const uint32_t n_cascades = cfg_u[0]; const int rgb_act = cfg_u[1], den_act = cfg_u[2], mode = cfg_u[3];
	for (uint32_t i = 0; i < n_rays; ++i) {
		RayPayload& p = payloads[i];
		if (!p.alive) continue;
		float* out = rgba + (size_t)i * 4;
		float l[4] = {out[0], out[1], out[2], out[3]};
		uint32_t j = 0;
		for (; j < p.n_steps; ++j) {
			const Coordinate& c = ((const Coordinate*)coords)[(size_t)i * n_steps + j];
			const uint16_t* o = net_out + ((size_t)i * n_steps + j) * padded_width;
			float dt = unwarp_dt(c.dt, min_step, n_cascades);
			float T = 1.f - l[3];
			float alpha = 1.f - expf(-net_to_density(h16from(o[3]), den_act) * dt);
			float w = alpha * T;
			float rgb[3] = {net_to_rgb(h16from(o[0]), rgb_act), net_to_rgb(h16from(o[1]), rgb_act), net_to_rgb(h16from(o[2]), rgb_act)};
			float diag = amax - amin;
			float pos[3] = {amin + c.px * diag, amin + c.py * diag, amin + c.pz * diag};
			if (mode == 3) {        // Depth
				float z = (cfg_f[5] * (pos[0] - p.ox) + cfg_f[6] * (pos[1] - p.oy) + cfg_f[7] * (pos[2] - p.oz)) * depth_scale;
				rgb[0] = rgb[1] = rgb[2] = z;
			} else if (mode == 4) { // Distance
				float dx = pos[0] - p.ox, dy = pos[1] - p.oy, dz = pos[2] - p.oz;
				float z = sqrtf(dx * dx + dy * dy + dz * dz) * depth_scale;
				rgb[0] = rgb[1] = rgb[2] = z;
			} else if (mode == 5) { // Stepsize
				rgb[0] = rgb[1] = rgb[2] = c.dt;
			} else if (mode == 0) { // AO
				rgb[0] = rgb[1] = rgb[2] = alpha;
			}
			for (int k = 0; k < 3; ++k) l[k] += rgb[k] * w;
			l[3] += w;
			if (l[3] > (1.0f - T_thresh)) break;
		}
		if (j < n_steps) { p.alive = 0; p.n_steps = j + n_steps_offset; }
		for (int k = 0; k < 4; ++k) out[k] = l[k];
	}
}

// frame[idx] = rgba + frame[idx]*(1-a), sRGB->linear when the model trains in sRGB and mode is Shade
void orc_shade(uint32_t n_hit, const float* rgba, const void* payloads_, float* frame, int render_mode, int train_in_linear) {
	const RayPayload* payloads = (const RayPayload*)payloads_;
	for (uint32_t i = 0; i < n_hit; ++i) {
		float t[4] = {rgba[i * 4], rgba[i * 4 + 1], rgba[i * 4 + 2], rgba[i * 4 + 3]};
		if (render_mode == 6) { float c = (float)payloads[i].n_steps / 128; t[0] = t[1] = t[2] = c; t[3] = 1.f; }
		if (!train_in_linear && render_mode == 1) for (int k = 0; k < 3; ++k) t[k] = srgb_to_linear(t[k]);
		float* f = frame + (size_t)payloads[i].idx * 4;
		for (int k = 0; k < 4; ++k) f[k] = t[k] + f[k] * (1.0f - t[3]);
	}
}

void orc_accumulate(uint32_t n_pix, const float* frame, float* accum, float sample_count, int color_space) {
	for (uint32_t i = 0; i < n_pix; ++i) {
		float c[4] = {frame[i * 4], frame[i * 4 + 1], frame[i * 4 + 2], frame[i * 4 + 3]};
		float* t = accum + (size_t)i * 4;
		if (color_space == 1) for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb(c[k]);
		for (int k = 0; k < 4; ++k) t[k] = (t[k] * sample_count + c[k]) / (sample_count + 1);
	}
}

// Identity tonemap curve only (the default; other curves are "next" rows)
void orc_tonemap(uint32_t n_pix, float exposure, const float* bg_rgba, const float* accum, int color_space, int out_color_space, float* out) {
	for (uint32_t i = 0; i < n_pix; ++i) {
		float c[4] = {accum[i * 4], accum[i * 4 + 1], accum[i * 4 + 2], accum[i * 4 + 3]};
		float scale = powf(2.0f, exposure);
		float bg[3] = {bg_rgba[0], bg_rgba[1], bg_rgba[2]};
		for (int k = 0; k < 3; ++k) c[k] *= scale;
		if (color_space != 1) for (int k = 0; k < 3; ++k) bg[k] = srgb_to_linear(bg[k]);
		float w = (1 - c[3]) * bg_rgba[3];
		for (int k = 0; k < 3; ++k) c[k] += bg[k] * w;
		c[3] += w;
		if (color_space == 1) for (int k = 0; k < 3; ++k) c[k] = srgb_to_linear(c[k]);
		if (out_color_space == 1) for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb(c[k]);
		for (int k = 0; k < 4; ++k) out[(size_t)i * 4 + k] = c[k];
	}
}

} // extern "C"
