// Testbed::render_frame / render_to_host (reference: internal/testbed.cu:479-605, internal/python_api.cu:22-46).
#include "testbed.h"
#include <algorithm>

namespace ngp {

void Testbed::render_frame(const Mat34& camera_matrix, CudaRenderBuffer& render_buffer, bool to_srgb) {
	if (!m_trainer) throw std::runtime_error("render: load_model_config or load_snapshot first");
	RenderingBuffer& rb = m_rendering_buffer;
	const int* resolution = render_buffer.resolution();
	const float focal_length[2] = {(float)resolution[rb.fov_axis] * rb.relative_focal_length[0], (float)resolution[rb.fov_axis] * rb.relative_focal_length[1]};
	// tile-sharded inference (BASELINE configs[3]): this rank marches the rays of rows [row_begin, row_end) only; the frame buffer
	// keeps the full size and the band's pixels land at their own positions, so the caller assembles bands without any exchange
	const int row_begin = m_render_row_end < 0 ? 0 : std::max(0, std::min(m_render_row_begin, resolution[1]));
	const int row_end = m_render_row_end < 0 ? resolution[1] : std::max(row_begin, std::min(m_render_row_end, resolution[1]));
	const uint32_t n_pixels = (uint32_t)(resolution[0] * (row_end - row_begin));
	const uint32_t n_frame_pixels = (uint32_t)(resolution[0] * resolution[1]);
	const ERenderMode render_mode = rb.render_mode;
	const uint32_t padded_output_width = m_network->padded_output_width();

	render_buffer.set_rows(row_begin, row_end);
	render_buffer.clear_frame_buffer(m_stream);
	(void)n_frame_pixels;

	if (rb.device_loop) {
		// Rounds are decided on the device (ngp_render_round): compaction, the step-count decision, march, network and compositing of
		// ROUNDS_PER_READBACK rounds are enqueued back to back, then the live-ray count comes back once.  Rounds enqueued after the
		// last ray died find a zero count and do nothing.  The launch bounds shrink with every read-back.
		// The rows are split into up to two bands that are marched CONCURRENTLY on two streams: the march and the compositing of a
		// round are latency bound and leave most of an SM idle, the network kernels in between are throughput bound, so one band's
		// march runs under the other band's encoder.  (Measured: 15.4 -> 14.9 ms per 1080p frame -- less than the idle slots
		// suggest: the march waits on the same L1TEX path the encoder saturates.  Staggering the bands by one march made no difference.)  A band is bit-identical to the same rows of a single-pass frame (the ray, its
		// jitter seed and its scatter target use the global pixel index; the grouping of a ray's samples into rounds does not matter).
		const uint32_t tail_steps = std::min<uint32_t>(std::max(rb.TAIL_STEPS, rb.MAX_STEPS_INBETWEEN_COMPACTION), 64u);
		const uint32_t per_batch = std::max(rb.ROUNDS_PER_READBACK, 1u);
		const int n_rows = row_end - row_begin;
		// (Normals needs the network's training scratch, which exists once: one band)
		const int n_bands = (rb.CONCURRENT_BANDS >= 2 && n_rows >= 2 && n_pixels >= rb.CONCURRENT_BANDS_MIN_PIXELS && render_mode != ERenderMode::Normals) ? 2 : 1;
		struct BandRun { int r0, r1; uint32_t n_pixels, bound, round; cudaStream_t stream; };
		BandRun run[2];
		if (n_bands > 1) NGP_CUDA_CHECK(cudaEventRecord(m_ev_frame_cleared, m_stream));
		for (int b = 0; b < n_bands; ++b) {
			BandRun& u = run[b];
			auto& bd = rb.band[b];
			u.r0 = row_begin + (int)((int64_t)n_rows * b / n_bands); u.r1 = row_begin + (int)((int64_t)n_rows * (b + 1) / n_bands);
			u.n_pixels = (uint32_t)(resolution[0] * (u.r1 - u.r0)); u.bound = u.n_pixels; u.round = 0;
			u.stream = n_bands > 1 ? m_band_stream[b] : m_stream;
			if (n_bands > 1) NGP_CUDA_CHECK(cudaStreamWaitEvent(u.stream, m_ev_frame_cleared, 0));
			const uint32_t n_elements = next_multiple(std::max(u.n_pixels, 1u), BATCH_SIZE_MULTIPLE);
			bd.rays[0].enlarge(n_elements); bd.rays[1].enlarge(n_elements); bd.rays_hit.enlarge(n_elements);
			m_sampler->generate_rays_from_camera_matrix(resolution, focal_length, rb.principal_point, camera_matrix, render_buffer.spp(), bd.rays[0].payload, u.stream,
			                                            u.r0, u.r1, m_train_envmap ? m_envmap.params_inference() : nullptr, m_envmap.resolution,
			                                            m_train_envmap ? render_buffer.frame_buffer() : nullptr);      // testbed.cu:506-512
			NGP_CUDA_CHECK(cudaMemsetAsync(bd.rays[0].rgba.data(), 0, (size_t)std::max(u.n_pixels, 1u) * 4 * sizeof(float), u.stream));
			NGP_CUDA_CHECK(cudaMemsetAsync(bd.n_hit_counter.data(), 0, sizeof(uint32_t), u.stream));
			NGP_CUDA_CHECK(cudaMemsetAsync(bd.round_ctl.data(), 0, sizeof(uint32_t) * NGP_RENDER_CTL_WORDS, u.stream));
		}
		rb.rounds_last_frame = rb.readbacks_last_frame = 0;
		while (true) {
			bool any = false;
			for (int b = 0; b < n_bands; ++b) {
				BandRun& u = run[b];
				if (u.bound == 0) continue;
				any = true;
				auto& bd = rb.band[b];
				uint32_t* ctl = bd.round_ctl.data();
				// samples of a round: n_alive * clamp(n_pixels / n_alive, MIN, MAX) <= max(min(n_pixels, bound * MAX), bound * MIN); the tail
				// rule stays below a quarter of the band
				const uint64_t batch = std::max(u.n_pixels, rb.MIN_BATCH_SAMPLES);
				const uint64_t by_rule = std::max<uint64_t>(std::min<uint64_t>(batch, (uint64_t)u.bound * rb.MAX_STEPS_INBETWEEN_COMPACTION), (uint64_t)u.bound * rb.MIN_STEPS_INBETWEEN_COMPACTION);
				const uint64_t by_tail = tail_steps > rb.MAX_STEPS_INBETWEEN_COMPACTION ? std::min<uint64_t>(batch / 4, (uint64_t)u.bound * tail_steps) : 0;
				const uint32_t sample_bound = next_multiple((uint32_t)std::max<uint64_t>(std::max(by_rule, by_tail), 1), BATCH_SIZE_MULTIPLE);
				bd.coords.enlarge(sample_bound);
				bd.rgbsigma.enlarge((size_t)sample_bound * padded_output_width);
				for (uint32_t k = 0; k < per_batch; ++k, ++u.round) {
					const uint32_t round = u.round;
					RaysSoA& rays_previous = bd.rays[round % 2];
					RaysSoA& rays_current = bd.rays[(round + 1) % 2];
					// only the frame's first round compacts in a launch of its own (the rays the set-up left alive); afterwards the
					// compositing of round r - 1 has already sorted its rays into this round's buffer and the hit list
					if (round == 0 || !rb.fused_compaction)
						api_check(ngp_compact_rays(u.stream, u.bound, round == 0 ? nullptr : ctl + (round & 1), rays_previous.payload.data(), rays_previous.rgba.data(),
						                           rays_current.payload.data(), rays_current.rgba.data(), bd.rays_hit.payload.data(), bd.rays_hit.rgba.data(),
						                           ctl + ((round + 1) & 1), bd.n_hit_counter.data()));
					api_check(ngp_render_round(u.stream, ctl, round, u.n_pixels, rb.MIN_STEPS_INBETWEEN_COMPACTION, rb.MAX_STEPS_INBETWEEN_COMPACTION, tail_steps, MARCH_ITER, rb.MIN_BATCH_SAMPLES));
					m_sampler->generate_coords_from_rays_dev(u.bound, ctl, round, rays_current.payload, bd.coords, u.stream);
					// rendering uses the EMA weights (use_inference_matrices = true, testbed.cu:556)
					m_network->inference_mixed_precision(u.stream, bd.coords.data(), sample_bound, ctl + 3, bd.rgbsigma.data(), true, n_bands > 1 ? &bd.workspace : nullptr);
					if (render_mode == ERenderMode::Normals) m_network->input_gradient(u.stream, 3, bd.coords.data(), sample_bound, ctl + 3);   // testbed.cu:558-560
					m_render->volume_rendering_dev(u.bound, ctl, round, rays_current, bd.coords, bd.rgbsigma, padded_output_width, m_density_activation, m_rgb_activation,
					                               render_mode, m_sampler->aabb, m_sampler->n_cascades, m_sampler->min_cone_stepsize, camera_matrix, 1.f / m_scene_scale, u.stream,
					                               rb.fused_compaction ? &rays_previous : nullptr, rb.fused_compaction ? &bd.rays_hit : nullptr,
					                               rb.fused_compaction ? bd.n_hit_counter.data() : nullptr);
				}
				// live rays after the batch's last compaction (some may have been finished by its compositing: the next compaction sorts
				// that out), still an upper bound for everything that follows
				// (with the fused compaction the newest count is the one the last compositing left for the next round)
				NGP_CUDA_CHECK(cudaMemcpyAsync(m_host_counter + 8 * b, ctl + ((u.round + (rb.fused_compaction ? 1u : 0u)) & 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, u.stream));
				NGP_CUDA_CHECK(cudaMemcpyAsync(m_host_counter + 8 * b + 1, ctl + 4, sizeof(uint32_t), cudaMemcpyDeviceToHost, u.stream));
			}
			if (!any) break;
			++rb.readbacks_last_frame;
			for (int b = 0; b < n_bands; ++b) {
				BandRun& u = run[b];
				if (u.bound == 0) continue;
				NGP_CUDA_CHECK(cudaStreamSynchronize(u.stream));
				u.bound = m_host_counter[8 * b + 1] >= MARCH_ITER ? 0 : m_host_counter[8 * b];
			}
		}
		for (int b = 0; b < n_bands; ++b) {
			auto& bd = rb.band[b];
			rb.rounds_last_frame = std::max(rb.rounds_last_frame, run[b].round);
			// the hit count stays on the device: no read-back before the shade
			api_check(ngp_shade(run[b].stream, std::max(run[b].n_pixels, 1u), bd.n_hit_counter.data(), bd.rays_hit.rgba.data(), bd.rays_hit.payload.data(),
			                    render_buffer.frame_buffer(), (int)render_mode, m_render->train_in_linear_color ? 1 : 0));
			if (n_bands > 1) {
				NGP_CUDA_CHECK(cudaEventRecord(m_ev_band_done[b], run[b].stream));
				NGP_CUDA_CHECK(cudaStreamWaitEvent(m_stream, m_ev_band_done[b], 0));
			}
		}
	} else {
	{
		const uint32_t n_elements = next_multiple(std::max(n_pixels, 1u), BATCH_SIZE_MULTIPLE);
		rb.rays[0].enlarge(n_elements); rb.rays[1].enlarge(n_elements); rb.rays_hit.enlarge(n_elements);
		rb.coords.enlarge((size_t)n_elements * rb.MAX_STEPS_INBETWEEN_COMPACTION);
		rb.rgbsigma.enlarge((size_t)n_elements * rb.MAX_STEPS_INBETWEEN_COMPACTION * padded_output_width);
	}
	m_sampler->generate_rays_from_camera_matrix(resolution, focal_length, rb.principal_point, camera_matrix, render_buffer.spp(), rb.rays[0].payload, m_stream,
	                                            row_begin, row_end, m_train_envmap ? m_envmap.params_inference() : nullptr, m_envmap.resolution,
	                                            m_train_envmap ? render_buffer.frame_buffer() : nullptr);      // testbed.cu:506-512
	NGP_CUDA_CHECK(cudaMemsetAsync(rb.rays[0].rgba.data(), 0, (size_t)std::max(n_pixels, 1u) * 4 * sizeof(float), m_stream));
	NGP_CUDA_CHECK(cudaMemsetAsync(rb.n_hit_counter.data(), 0, sizeof(uint32_t), m_stream));

	uint32_t i_steps_offset = 1;
	uint32_t double_buffer_index = 0;
	uint32_t n_alive = n_pixels;
	rb.rounds_last_frame = rb.readbacks_last_frame = 0;
	while (n_alive > 0 && i_steps_offset < MARCH_ITER) {
		RaysSoA& rays_previous = rb.rays[double_buffer_index % 2];
		RaysSoA& rays_current = rb.rays[(double_buffer_index + 1) % 2];
		++double_buffer_index;

		NGP_CUDA_CHECK(cudaMemsetAsync(rb.n_alive_counter.data(), 0, sizeof(uint32_t), m_stream));
		api_check(ngp_compact_rays(m_stream, n_alive, nullptr, rays_previous.payload.data(), rays_previous.rgba.data(), rays_current.payload.data(), rays_current.rgba.data(),
		                           rb.rays_hit.payload.data(), rb.rays_hit.rgba.data(), rb.n_alive_counter.data(), rb.n_hit_counter.data()));
		NGP_CUDA_CHECK(cudaMemcpyAsync(m_host_counter, rb.n_alive_counter.data(), sizeof(uint32_t), cudaMemcpyDeviceToHost, m_stream));
		NGP_CUDA_CHECK(cudaStreamSynchronize(m_stream));
		++rb.readbacks_last_frame;
		n_alive = m_host_counter[0];
		if (n_alive == 0) break;
		++rb.rounds_last_frame;

		const uint32_t n_steps = std::min(std::max(n_pixels / n_alive, rb.MIN_STEPS_INBETWEEN_COMPACTION), rb.MAX_STEPS_INBETWEEN_COMPACTION);
		m_sampler->generate_coords_from_rays(n_alive, nullptr, n_steps, rays_current.payload, rb.coords, m_stream);
		// rendering uses the EMA weights (use_inference_matrices = true, testbed.cu:556)
		m_network->inference_mixed_precision(m_stream, rb.coords.data(), n_alive * n_steps, nullptr, rb.rgbsigma.data(), true);
		if (render_mode == ERenderMode::Normals) m_network->input_gradient(m_stream, 3, rb.coords.data(), n_alive * n_steps);   // testbed.cu:558-560
		m_render->volume_rendering(n_alive, nullptr, n_steps, i_steps_offset, rays_current, rb.coords, rb.rgbsigma, padded_output_width, m_density_activation,
		                           m_rgb_activation, render_mode, m_sampler->aabb, m_sampler->n_cascades, m_sampler->min_cone_stepsize, camera_matrix,
		                           1.f / m_scene_scale, m_stream);
		i_steps_offset += n_steps;
	}

	NGP_CUDA_CHECK(cudaMemcpyAsync(m_host_counter + 1, rb.n_hit_counter.data(), sizeof(uint32_t), cudaMemcpyDeviceToHost, m_stream));
	NGP_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	const uint32_t n_hit = m_host_counter[1];
	api_check(ngp_shade(m_stream, n_hit, nullptr, rb.rays_hit.rgba.data(), rb.rays_hit.payload.data(), render_buffer.frame_buffer(), (int)render_mode,
	                    m_render->train_in_linear_color ? 1 : 0));
	}

	render_buffer.set_color_space(m_color_space);
	render_buffer.set_tonemap_curve(rb.tonemap_curve);
	render_buffer.accumulate(m_stream);
	render_buffer.tonemap(rb.exposure, m_background_color, to_srgb ? EColorSpace::SRGB : EColorSpace::Linear, m_stream);
	NGP_CUDA_CHECK(cudaStreamSynchronize(m_stream));
}

void Testbed::render_to_host(int width, int height, int spp, bool to_srgb, float* out) {
	RenderingBuffer& rb = m_rendering_buffer;
	rb.render_buffer.resize(width, height);
	rb.render_buffer.reset_accumulation();
	for (int i = 0; i < spp; ++i) render_frame(rb.camera_matrix, rb.render_buffer, to_srgb);
	if (m_render_row_end < 0) {
		NGP_CUDA_CHECK(cudaMemcpy(out, rb.render_buffer.surface(), (size_t)width * height * 4 * sizeof(float), cudaMemcpyDeviceToHost));
		return;
	}
	// a band: only its rows were rendered and only they are copied; the rest of the caller's image reads zero
	const size_t first = rb.render_buffer.first_pixel() * 4, n = (size_t)rb.render_buffer.n_band_pixels() * 4, total = (size_t)width * height * 4;
	std::fill(out, out + first, 0.f);
	std::fill(out + first + n, out + total, 0.f);
	if (n) NGP_CUDA_CHECK(cudaMemcpy(out + first, rb.render_buffer.surface() + first, n * sizeof(float), cudaMemcpyDeviceToHost));
}

// the band alone, [row_end - row_begin][width][4]: what a rank of a tile-sharded render hands on (no full-frame traffic at all)
void Testbed::render_band_to_host(int width, int height, int spp, bool to_srgb, float* out_band) {
	RenderingBuffer& rb = m_rendering_buffer;
	rb.render_buffer.resize(width, height);
	rb.render_buffer.reset_accumulation();
	for (int i = 0; i < spp; ++i) render_frame(rb.camera_matrix, rb.render_buffer, to_srgb);
	const size_t first = rb.render_buffer.first_pixel() * 4, n = (size_t)rb.render_buffer.n_band_pixels() * 4;
	if (n) NGP_CUDA_CHECK(cudaMemcpy(out_band, rb.render_buffer.surface() + first, n * sizeof(float), cudaMemcpyDeviceToHost));
}

float Testbed::render_frames_ms(int width, int height, int spp, bool to_srgb) {
	RenderingBuffer& rb = m_rendering_buffer;
	rb.render_buffer.resize(width, height);
	rb.render_buffer.reset_accumulation();
	NGP_CUDA_CHECK(cudaDeviceSynchronize());
	const auto t0 = std::chrono::steady_clock::now();
	for (int i = 0; i < spp; ++i) render_frame(rb.camera_matrix, rb.render_buffer, to_srgb);
	NGP_CUDA_CHECK(cudaDeviceSynchronize());
	return std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

void Testbed::set_camera_to_training_view(uint32_t i) {
	const NeRFSynthetic& d = m_training_buffer.dataset;
	if (i >= d.n_images) throw std::runtime_error("set_camera_to_training_view: no such view");
	m_rendering_buffer.camera_matrix = d.xforms_host[i];
	const int axis = (int)m_rendering_buffer.fov_axis;
	m_rendering_buffer.relative_focal_length[0] = d.focal_length[0] / (float)d.image_resolution[axis];
	m_rendering_buffer.relative_focal_length[1] = d.focal_length[1] / (float)d.image_resolution[axis];
}

void Testbed::training_image_to_host(uint32_t i, float* out) const {
	const NeRFSynthetic& d = m_training_buffer.dataset;
	if (i >= d.n_images) throw std::runtime_error("training_image_to_host: no such view");
	const size_t n = (size_t)d.image_resolution[0] * d.image_resolution[1] * 4;
	std::vector<uint16_t> h(n);
	NGP_CUDA_CHECK(cudaMemcpy(h.data(), d.images.data() + n * i, n * 2, cudaMemcpyDeviceToHost));
	for (size_t k = 0; k < n; ++k) {
		const uint32_t s = (h[k] >> 15) & 1, e = (h[k] >> 10) & 0x1f, m = h[k] & 0x3ff;
		float v = e == 0 ? std::ldexp((float)m, -24) : e == 31 ? (m ? NAN : INFINITY) : std::ldexp((float)(m | 0x400), (int)e - 25);
		out[k] = s ? -v : v;
	}
}

} // namespace ngp
