// TEST INFRASTRUCTURE ONLY -- a small driver around the UNMODIFIED reference (Testbed and its module classes, compiled from
// /root/reference by oracle/build_ref.sh).  It feeds the reference the same synthetic dataset our path gets, runs the
// reference's own Testbed::train_prep_nerf / train_nerf / train, and either dumps the buffers the parity tests compare
// against (dump) or prints the training throughput (bench).  Nothing of the product links or includes this file.
//
//   ref_harness dump  <dir> [batch] [n_rays]   reads <dir>/{config.json,meta.txt,xforms.f32,images.f16}, writes <dir>/ref_*.bin
//   ref_harness bench <dir> <calls> <warmup>   train(16, batch) loop as experiment/run.py:71 drives it; one JSON line
//   ref_harness trainrender <dir> <calls> <W> <H> <spp> <fov>   train, then render <dir>/test_xforms.f32 -> ref_render_<i>.f32
//   ref_harness benchsteps <dir> <steps> <warmup_steps> [batch] [psnr_res] [W] [H] [fov]   the same loop for ANY step counts (calls of 16 plus one remainder
//                                              call); with <dir>/test_xforms.f32 the poses are then rendered at psnr_res^2 (written out) and timed at W x H
//   ref_harness snapshot <dir> <calls> <W> <H> <fov> <specs> [batch]   train, Testbed::save_snapshot (with and without optimizer state), render test_xforms.f32
//   ref_harness loadrender <dir> <file.msgpack> <W> <H> <fov> <specs> [train_calls] [batch]   Testbed::load_snapshot of ANY file (e.g. one written by
//                                              pyngp), render the poses (timed), optionally keep training from the loaded state
//   ref_harness renderdump <dir> <file.msgpack> <W> <H> <fov> <mode> <rounds>   Testbed::render_frame replayed stage by stage on pose 0, buffers dumped
//   ref_harness loaddata <dir> <transforms.json> <scale>   Testbed::load_training_data (stb_image PNG decode, from_rgba32, spec_opengl_to_opencv) dumped
//      <specs> = comma-separated mode:curve:spp:srgb[:exposure] (ERenderMode / ETonemapCurve ints), one output image per spec and pose
//   ref_harness c1 <dir> <iters> [log2_T]      BASELINE configs[0]: the reference's own GridEncoding (tcnn::create_encoding) on <dir>/c1_xyz.f32
//                                              ([N][3]), c1_table.f16, c1_dy.f16 ([N][32]): times encode / backward, writes ref_c1_enc.f16, ref_c1_grad.f16
#include "internal/testbed.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

using namespace Eigen;

template <typename T>
static std::vector<T> read_file(const std::string& path) {
	std::ifstream f(path, std::ios::binary | std::ios::ate);
	if (!f) throw std::runtime_error("cannot open " + path);
	const size_t bytes = (size_t)f.tellg();
	std::vector<T> v(bytes / sizeof(T));
	f.seekg(0);
	f.read((char*)v.data(), (std::streamsize)(v.size() * sizeof(T)));
	return v;
}

static void write_device(const std::string& path, const void* dptr, size_t bytes) {
	std::vector<char> h(bytes);
	if (bytes) CUDA_CHECK_THROW(cudaMemcpy(h.data(), dptr, bytes, cudaMemcpyDeviceToHost));
	std::ofstream f(path, std::ios::binary);
	f.write(h.data(), (std::streamsize)bytes);
}

static void write_host(const std::string& path, const void* p, size_t bytes) {
	std::ofstream f(path, std::ios::binary);
	f.write((const char*)p, (std::streamsize)bytes);
}

// what Testbed::load_training_data does (internal/testbed.cu:95-125) with the decoded images handed over directly
static void load_dataset(Testbed& tb, const std::string& dir) {
	std::ifstream meta(dir + "/meta.txt");
	size_t n_images; int W, H; float fx, fy, scale;
	meta >> n_images >> W >> H >> fx >> fy >> scale;
	if (!meta) throw std::runtime_error("bad meta.txt");
	auto xf = read_file<float>(dir + "/xforms.f32");            // [n][12] column-major, already in the NGP frame
	auto img = read_file<uint16_t>(dir + "/images.f16");        // [n][H][W][4] linear premultiplied fp16
	if (xf.size() != n_images * 12 || img.size() != n_images * (size_t)W * H * 4) throw std::runtime_error("dataset size mismatch");
	NeRFSynthetic& d = tb.m_training_buffer.dataset;
	d.scale = scale; d.offset = Vector3f::Constant(0.5f);
	d.n_images = n_images;
	d.image_resolution = Vector2i(W, H);
	d.focal_length = Vector2f(fx, fy);
	d.principal_point = Vector2f::Constant(0.5f);
	d.xforms.resize(n_images);
	CUDA_CHECK_THROW(cudaMemcpy(d.xforms.data(), xf.data(), xf.size() * sizeof(float), cudaMemcpyHostToDevice));
	d.images.resize(img.size());
	CUDA_CHECK_THROW(cudaMemcpy(d.images.data(), img.data(), img.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
	tb.m_scene_scale = scale;
	tb.m_scene_offset = Vector3f::Constant(0.5f);
	tb.m_sampler->init((uint32_t)d.n_images, d.image_resolution, d.focal_length, d.principal_point, d.xforms, d.images, tb.m_rng, tb.m_stream);
	CUDA_CHECK_THROW(cudaStreamSynchronize(tb.m_stream));
}

static int run_dump(const std::string& dir, uint32_t batch, uint32_t n_rays_override) {
	Testbed tb;
	if (n_rays_override) tb.m_training_buffer.n_rays_per_batch = n_rays_override;
	tb.load_model_config(dir + "/config.json");
	const size_t n = tb.m_network->n_params();
	float* p32 = tb.m_trainer->params();
	const __half* p16 = (const __half*)(p32 + n);               // arena order of Trainer::initialize_params (trainer.h:79-85)
	const __half* grads = p16 + 2 * n;
	CUDA_CHECK_THROW(cudaDeviceSynchronize());
	write_device(dir + "/ref_params_init_f32.bin", p32, n * 4);  // what Trainer::initialize_params produced (seed 43)
	{
		// optional: start from given fp16 parameters instead (Trainer::set_params), so the numerics are exercised at O(1) values
		std::ifstream probe(dir + "/params_init.f16", std::ios::binary);
		if (probe.good()) {
			auto h = read_file<uint16_t>(dir + "/params_init.f16");
			if (h.size() != n) throw std::runtime_error("params_init.f16 has the wrong size");
			tb.m_trainer->set_params((const precision_t*)h.data(), n);
			CUDA_CHECK_THROW(cudaDeviceSynchronize());
		}
	}
	write_device(dir + "/ref_params0_f32.bin", p32, n * 4);
	const size_t n_env = tb.m_envmap.envmap->n_params();
	if (tb.m_train_envmap) {
		// start the environment map from given texels (Trainer::set_params), so its read and its gradient deposit are exercised
		std::ifstream probe(dir + "/envmap_init.f16", std::ios::binary);
		if (probe.good()) {
			auto h = read_file<uint16_t>(dir + "/envmap_init.f16");
			if (h.size() != n_env) throw std::runtime_error("envmap_init.f16 has the wrong size");
			tb.m_envmap.trainer->set_params((const precision_t*)h.data(), n_env);
			CUDA_CHECK_THROW(cudaDeviceSynchronize());
		}
	}

	load_dataset(tb, dir);
	OccupancySampler& s = *tb.m_sampler;
	write_device(dir + "/ref_grid0_f32.bin", s.density_grid.data(), (size_t)s.n_total_elements * 4);
	write_device(dir + "/ref_bitfield0_u8.bin", s.density_grid_bitfield.data(), s.density_grid_bitfield.size());

	// occupancy refresh (internal/testbed.cu:193-229); the sampler's own generator going in is recorded
	const uint64_t sampler_rng[2] = {s.rng.state, s.rng.inc};
	tb.train_prep_nerf(batch, 1, tb.m_stream);
	CUDA_CHECK_THROW(cudaStreamSynchronize(tb.m_stream));
	auto& b = tb.m_training_buffer;
	write_device(dir + "/ref_grid_positions_f32.bin", b.grid_positions.data(), (size_t)s.n_total_elements * sizeof(Position));
	write_device(dir + "/ref_grid_indices_u32.bin", b.grid_indices.data(), (size_t)s.n_total_elements * 4);
	write_device(dir + "/ref_grid_density_f16.bin", b.rgbsigma.data(), (size_t)s.n_total_elements * tb.m_network->padded_density_output_width() * 2);
	write_device(dir + "/ref_grid1_f32.bin", s.density_grid.data(), (size_t)s.n_total_elements * 4);
	write_device(dir + "/ref_grid_mean_f32.bin", s.density_grid_mean.data(), 4);
	write_device(dir + "/ref_bitfield1_u8.bin", s.density_grid_bitfield.data(), s.density_grid_bitfield.size());

	// One training step, stage by stage, through the reference's own module methods in the order Testbed::train_nerf_step
	// issues them (internal/testbed.cu:301-418), with the buffers dumped between stages.  (bench mode below calls the
	// reference's Testbed::train itself.)
	cudaStream_t stream = tb.m_stream;
	const uint64_t rng_in[2] = {tb.m_rng.state, tb.m_rng.inc};
	const uint32_t n_rays = b.n_rays_per_batch, max_samples = batch * 16, w = tb.m_network->padded_output_width();
	b.rays.enlarge(n_rays); b.rays_rgba.enlarge(n_rays); b.ray_counter.enlarge(1); b.ray_coord_indices.enlarge(n_rays * 2);
	b.coords.enlarge(max_samples); b.coords_compacted.enlarge(batch);
	b.rgbsigma.enlarge((size_t)max_samples * w); b.dloss_drgbsigma.enlarge((size_t)batch * w);
	b.n_coords_counter.enlarge(1); b.n_compacted_coords_counter.enlarge(1); b.loss.enlarge(n_rays);
	CUDA_CHECK_THROW(cudaMemsetAsync(b.ray_counter.data(), 0, 4, stream));
	CUDA_CHECK_THROW(cudaMemsetAsync(b.n_coords_counter.data(), 0, 4, stream));
	CUDA_CHECK_THROW(cudaMemsetAsync(b.n_compacted_coords_counter.data(), 0, 4, stream));
	CUDA_CHECK_THROW(cudaMemsetAsync(b.loss.data(), 0, (size_t)n_rays * 4, stream));
	CUDA_CHECK_THROW(cudaMemsetAsync(b.coords_compacted.data(), 0, (size_t)batch * sizeof(Coordinate), stream));
	CUDA_CHECK_THROW(cudaMemsetAsync(b.dloss_drgbsigma.data(), 0, (size_t)batch * w * 2, stream));
	const uint32_t max_inference = tcnn::next_multiple(max_samples, BATCH_SIZE_MULTIPLE);

	precision_t* envmap_gradient = tb.m_train_envmap ? tb.m_envmap.envmap->gradients() : nullptr;     // testbed.cu:236-239
	if (envmap_gradient) CUDA_CHECK_THROW(cudaMemsetAsync(envmap_gradient, 0, sizeof(precision_t) * n_env, stream));
	tb.m_sampler->sample_coords_from_dataset(n_rays, 0, max_inference, b.dataset, tb.m_train_envmap, b.rays, b.rays_rgba, b.ray_counter.data(), b.ray_coord_indices, b.coords,
	                                         b.n_coords_counter.data(), tb.m_rng, stream);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	uint32_t n_kept = 0, n_marched = 0, n_comp = 0;
	CUDA_CHECK_THROW(cudaMemcpy(&n_kept, b.ray_counter.data(), 4, cudaMemcpyDeviceToHost));
	CUDA_CHECK_THROW(cudaMemcpy(&n_marched, b.n_coords_counter.data(), 4, cudaMemcpyDeviceToHost));
	const uint32_t n_used = std::min(n_marched, max_inference);
	write_device(dir + "/ref_rays_f32.bin", b.rays.data(), (size_t)n_kept * sizeof(Ray));
	write_device(dir + "/ref_rays_rgba_f32.bin", b.rays_rgba.data(), (size_t)n_kept * 16);
	write_device(dir + "/ref_numsteps_u32.bin", b.ray_coord_indices.data(), (size_t)n_kept * 8);
	write_device(dir + "/ref_coords_f32.bin", b.coords.data(), (size_t)n_used * sizeof(Coordinate));

	tcnn::GPUMatrix<float> coords_matrix((float*)b.coords.data(), sizeof(Coordinate) / sizeof(float), max_inference);
	tcnn::GPUMatrix<precision_t> rgbsigma_matrix(b.rgbsigma.data(), w, max_inference);
	tb.m_network->inference_mixed_precision(stream, coords_matrix, rgbsigma_matrix, false);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	write_device(dir + "/ref_rgbsigma_f16.bin", b.rgbsigma.data(), (size_t)n_used * w * 2);

	tb.m_render->volume_rendering_with_loss_and_gradient(
		n_rays, batch, b.rays, b.rays_rgba, b.ray_counter, b.ray_coord_indices, b.coords, b.coords_compacted, b.n_compacted_coords_counter.data(),
		b.rgbsigma, b.dloss_drgbsigma, w, tb.m_rgb_activation, tb.m_density_activation, b.loss.data(), tb.m_loss_type, LOSS_SCALE,
		tb.m_sampler->aabb, tb.m_sampler->n_cascades, tb.m_sampler->min_cone_stepsize, tb.m_sampler->density_grid_mean, tb.m_sampler->min_optical_thickness,
		tb.m_envmap.envmap->params(), envmap_gradient, tb.m_envmap.resolution, tb.m_envmap.loss_type,
		tb.m_background_color.head<3>(), tb.m_color_space, tb.m_rng, stream);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	CUDA_CHECK_THROW(cudaMemcpy(&n_comp, b.n_compacted_coords_counter.data(), 4, cudaMemcpyDeviceToHost));
	write_device(dir + "/ref_numsteps_compacted_u32.bin", b.ray_coord_indices.data(), (size_t)n_kept * 8);
	write_device(dir + "/ref_coords_compacted_f32.bin", b.coords_compacted.data(), (size_t)batch * sizeof(Coordinate));
	write_device(dir + "/ref_dloss_f16.bin", b.dloss_drgbsigma.data(), (size_t)batch * w * 2);
	write_device(dir + "/ref_loss_f32.bin", b.loss.data(), (size_t)n_rays * 4);
	if (envmap_gradient) {
		write_device(dir + "/ref_envmap_grad_f16.bin", envmap_gradient, n_env * 2);
		write_device(dir + "/ref_envmap_params0_f16.bin", tb.m_envmap.envmap->params(), n_env * 2);
	}

	tcnn::fill_rollover_and_rescale<precision_t><<<tcnn::n_blocks_linear(batch * w), tcnn::n_threads_linear, 0, stream>>>(
		batch, w, b.n_compacted_coords_counter.data(), b.dloss_drgbsigma.data());
	tcnn::fill_rollover<Coordinate><<<tcnn::n_blocks_linear(batch), tcnn::n_threads_linear, 0, stream>>>(batch, 1, b.n_compacted_coords_counter.data(), b.coords_compacted.data());
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	write_device(dir + "/ref_coords_rolled_f32.bin", b.coords_compacted.data(), (size_t)batch * sizeof(Coordinate));
	write_device(dir + "/ref_dloss_rolled_f16.bin", b.dloss_drgbsigma.data(), (size_t)batch * w * 2);

	tcnn::GPUMatrix<float> cc_matrix((float*)b.coords_compacted.data(), sizeof(Coordinate) / sizeof(float), batch);
	tcnn::GPUMatrix<precision_t> out_matrix(b.rgbsigma.data(), w, batch);
	tcnn::GPUMatrix<precision_t> grad_matrix(b.dloss_drgbsigma.data(), w, batch);
	tb.m_network->forward(stream, cc_matrix, &out_matrix);
	tb.m_network->backward(stream, cc_matrix, out_matrix, grad_matrix);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	write_device(dir + "/ref_rgbsigma_compacted_f16.bin", b.rgbsigma.data(), (size_t)batch * w * 2);
	write_device(dir + "/ref_grads_f16.bin", grads, n * 2);

	tb.m_rng.advance();
	tb.m_trainer->optimizer_step(stream, LOSS_SCALE);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	write_device(dir + "/ref_params1_f32.bin", p32, n * 4);
	write_device(dir + "/ref_params1_f16.bin", p16, n * 2);
	if (envmap_gradient) {
		tb.m_envmap.trainer->optimizer_step(stream, LOSS_SCALE * 1.0f);     // testbed.cu:295-298 with n_training_steps = 1
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		write_device(dir + "/ref_envmap_params1_f32.bin", tb.m_envmap.envmap->params_full_precision(), n_env * 4);
		write_device(dir + "/ref_envmap_params1_f16.bin", tb.m_envmap.envmap->params(), n_env * 2);
		write_device(dir + "/ref_envmap_inference1_f16.bin", tb.m_envmap.envmap->params_inference(), n_env * 2);
	}
	{
		std::ofstream f(dir + "/ref_step.txt");
		f << "rng_state " << rng_in[0] << "\nrng_inc " << rng_in[1] << "\nsampler_rng_state " << sampler_rng[0] << "\nsampler_rng_inc " << sampler_rng[1]
		  << "\nn_rays " << n_rays << "\nn_kept " << n_kept << "\nn_marched " << n_marched << "\nn_compacted " << n_comp << "\nbatch " << batch
		  << "\npadded_output_width " << w << "\nn_params " << n << "\nn_total_elements " << s.n_total_elements
		  << "\nn_cascades " << s.n_cascades << "\nn_envmap_params " << (tb.m_train_envmap ? n_env : 0) << "\n";
	}
	std::cout << "dumped: n_rays " << n_rays << " kept " << n_kept << " marched " << n_marched << " compacted " << n_comp << std::endl;
	return 0;
}

static int run_bench(const std::string& dir, uint32_t calls, uint32_t warmup, uint32_t batch) {
	Testbed tb;
	tb.load_model_config(dir + "/config.json");
	load_dataset(tb, dir);
	auto& b = tb.m_training_buffer;
	for (uint32_t i = 0; i < warmup; ++i) tb.train(16, batch);
	CUDA_CHECK_THROW(cudaDeviceSynchronize());
	double ms = 0, samples = 0, marched = 0, rays = 0;
	const auto t0 = std::chrono::steady_clock::now();
	for (uint32_t i = 0; i < calls; ++i) {
		const uint32_t r = b.n_rays_per_batch;
		tb.train(16, batch);
		ms += b.training_ms + b.training_prep_ms;                 // the reference's own timers (internal/testbed.cu:170-190)
		for (uint32_t k = 0; k < 16; ++k) { samples += std::min(b.n_compacted_coords_counter_cpu[k], batch); marched += b.n_coords_counter_cpu[k]; }
		rays += 16.0 * r;
	}
	CUDA_CHECK_THROW(cudaDeviceSynchronize());
	const double wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
	const double steps = 16.0 * calls;
	printf("{\"impl\": \"reference-cuda-sm100\", \"samples_per_s\": %.1f, \"ms_per_step\": %.5f, \"wall_ms_per_step\": %.5f, \"steps\": %.0f, \"warmup_steps\": %u, "
	       "\"samples_per_step\": %.1f, \"marched_per_step\": %.1f, \"rays_per_step\": %.1f, \"loss\": %.6g}\n",
	       samples / (ms * 1e-3), ms / steps, wall_ms / steps, steps, warmup * 16, samples / steps, marched / steps, rays / steps, (double)b.loss_scalar);
	return 0;
}

// train n_calls x 16 steps, then render the poses of <dir>/test_xforms.f32 the way Testbed::render (internal/python_api.cu:22-46) does
static int run_trainrender(const std::string& dir, uint32_t calls, int W, int H, int spp, float fov_value, uint32_t batch) {
	Testbed tb;
	tb.load_model_config(dir + "/config.json");
	load_dataset(tb, dir);
	auto& b = tb.m_training_buffer;
	std::ofstream lf(dir + "/ref_loss_curve.txt");
	double ms = 0;
	for (uint32_t i = 0; i < calls; ++i) {
		tb.train(16, batch);
		ms += b.training_ms + b.training_prep_ms;
		lf << (i + 1) * 16 << " " << b.loss_scalar << " " << b.measured_batch_size << " " << b.n_rays_per_batch << "\n";
	}
	auto poses = read_file<float>(dir + "/test_xforms.f32");
	const size_t n_poses = poses.size() / 12;
	tb.set_fov(fov_value);
	auto& rb = tb.m_rendering_buffer;
	std::vector<float> image((size_t)W * H * 4);
	double render_ms = 0;
	for (size_t p = 0; p < n_poses; ++p) {
		Eigen::Matrix<float, 3, 4> cam;
		for (int k = 0; k < 12; ++k) cam.data()[k] = poses[p * 12 + k];          // column-major, already in the NGP frame
		rb.camera_matrix = cam;
		CUDA_CHECK_THROW(cudaDeviceSynchronize());
		const auto t0 = std::chrono::steady_clock::now();
		rb.render_buffer.resize({W, H});
		rb.render_buffer.reset_accumulation();
		for (int i = 0; i < spp; ++i) tb.render_frame(rb.camera_matrix, rb.render_buffer, true);
		CUDA_CHECK_THROW(cudaDeviceSynchronize());
		if (p > 0) render_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();   // pose 0 warms up
		CUDA_CHECK_THROW(cudaMemcpy2DFromArray(image.data(), (size_t)W * 4 * sizeof(float), rb.render_buffer.surface_provider().array(), 0, 0,
		                                       (size_t)W * 4 * sizeof(float), H, cudaMemcpyDeviceToHost));
		write_host(dir + "/ref_render_" + std::to_string(p) + ".f32", image.data(), image.size() * 4);
	}
	const double mrays = n_poses > 1 ? (double)W * H * spp * (n_poses - 1) / (render_ms * 1e-3) / 1e6 : 0.0;
	printf("{\"impl\": \"reference-cuda-sm100\", \"train_steps\": %u, \"train_ms_per_step\": %.5f, \"loss\": %.6g, \"render_mrays_per_s\": %.3f, \"render_ms_per_frame\": %.3f, "
	       "\"n_poses\": %zu, \"width\": %d, \"height\": %d, \"spp\": %d}\n",
	       calls * 16, ms / (16.0 * calls), (double)b.loss_scalar, mrays, n_poses > 1 ? render_ms / (n_poses - 1) : 0.0, n_poses, W, H, spp);
	return 0;
}


struct RenderSpec { int mode = 1, curve = 0, spp = 1, srgb = 1; float exposure = 0.f; };
static std::vector<RenderSpec> parse_specs(const std::string& text);
static double render_specs(Testbed& tb, const std::string& dir, const std::string& prefix, int W, int H, float fov_value, const std::vector<RenderSpec>& specs, double* n_timed_rays);

// the training loop for arbitrary step counts: calls of 16 steps plus one remainder call (Testbed::train takes any n); then, when
// <dir>/test_xforms.f32 exists, the held-out poses rendered linear at psnr_res^2 (-> ref_bench_render_0_<p>.f32, for PSNR) and timed at W x H
static int run_benchsteps(const std::string& dir, uint32_t steps, uint32_t warmup, uint32_t batch, int psnr_res, int W, int H, float fov_value) {
	Testbed tb;
	tb.load_model_config(dir + "/config.json");
	load_dataset(tb, dir);
	auto& b = tb.m_training_buffer;
	for (uint32_t done = 0; done < warmup;) { const uint32_t k = std::min(16u, warmup - done); tb.train(k, batch); done += k; }
	CUDA_CHECK_THROW(cudaDeviceSynchronize());
	double ms = 0, samples = 0, marched = 0, rays = 0;
	const auto t0 = std::chrono::steady_clock::now();
	for (uint32_t done = 0; done < steps;) {
		const uint32_t k = std::min(16u, steps - done), r = b.n_rays_per_batch;
		tb.train(k, batch);
		ms += b.training_ms + b.training_prep_ms;
		for (uint32_t j = 0; j < k; ++j) { samples += std::min(b.n_compacted_coords_counter_cpu[j], batch); marched += b.n_coords_counter_cpu[j]; }
		rays += (double)k * r;
		done += k;
	}
	CUDA_CHECK_THROW(cudaDeviceSynchronize());
	const double wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
	double render_mrays = 0, render_ms = 0;
	int n_frames = 0;
	{
		std::ifstream probe(dir + "/test_xforms.f32", std::ios::binary);
		if (probe.good() && psnr_res > 0) {
			double nrays = 0;
			render_specs(tb, dir, "ref_bench_render", psnr_res, psnr_res, fov_value, parse_specs("1:0:1:0"), &nrays);
			if (W > 0 && H > 0) {
				// the same poses twice at the benchmark resolution; the very first frame (allocation) is not timed
				const double sec = render_specs(tb, dir, "ref_bench_frame", W, H, fov_value, parse_specs("1:0:1:1,1:0:1:1"), &nrays);
				n_frames = (int)(nrays / ((double)W * H));
				if (sec > 0 && n_frames > 0) { render_mrays = nrays / sec / 1e6; render_ms = sec * 1e3 / n_frames; }
			}
		}
	}
	printf("{\"impl\": \"reference-cuda-sm100\", \"samples_per_s\": %.1f, \"ms_per_step\": %.5f, \"wall_ms_per_step\": %.5f, \"steps\": %u, \"warmup_steps\": %u, "
	       "\"samples_per_step\": %.1f, \"marched_per_step\": %.1f, \"rays_per_step\": %.1f, \"loss\": %.6g, \"i_step\": %u, "
	       "\"render_mrays_per_s\": %.3f, \"render_ms_per_frame\": %.3f, \"render_frames_timed\": %d, \"render_width\": %d, \"render_height\": %d}\n",
	       samples / (ms * 1e-3), ms / steps, wall_ms / steps, steps, warmup, samples / steps, marched / steps, rays / steps, (double)b.loss_scalar, b.i_step,
	       render_mrays, render_ms, n_frames, W, H);
	return 0;
}

static std::vector<RenderSpec> parse_specs(const std::string& text) {
	std::vector<RenderSpec> out;
	size_t pos = 0;
	while (pos <= text.size()) {
		const size_t end = std::min(text.find(',', pos), text.size());
		const std::string item = text.substr(pos, end - pos);
		if (!item.empty()) {
			RenderSpec r;
			float f[5] = {1, 0, 1, 1, 0};
			size_t q = 0;
			for (int k = 0; k < 5 && q <= item.size(); ++k) { const size_t e = std::min(item.find(':', q), item.size()); if (e > q) f[k] = (float)atof(item.substr(q, e - q).c_str()); q = e + 1; }
			r.mode = (int)f[0]; r.curve = (int)f[1]; r.spp = std::max(1, (int)f[2]); r.srgb = (int)f[3]; r.exposure = f[4];
			out.push_back(r);
		}
		pos = end + 1;
	}
	if (out.empty()) out.push_back(RenderSpec{});
	return out;
}

// Testbed::render (internal/python_api.cu:22-46) for every spec x pose of <dir>/test_xforms.f32; returns the seconds spent in the timed frames
static double render_specs(Testbed& tb, const std::string& dir, const std::string& prefix, int W, int H, float fov_value, const std::vector<RenderSpec>& specs, double* n_timed_rays) {
	auto poses = read_file<float>(dir + "/test_xforms.f32");
	const size_t n_poses = poses.size() / 12;
	tb.set_fov(fov_value);
	auto& rb = tb.m_rendering_buffer;
	std::vector<float> image((size_t)W * H * 4);
	double seconds = 0; *n_timed_rays = 0;
	for (size_t s = 0; s < specs.size(); ++s) {
		rb.render_mode = (ERenderMode)specs[s].mode;
		rb.tonemap_curve = (ETonemapCurve)specs[s].curve;
		rb.exposure = specs[s].exposure;
		for (size_t p = 0; p < n_poses; ++p) {
			Eigen::Matrix<float, 3, 4> cam;
			for (int k = 0; k < 12; ++k) cam.data()[k] = poses[p * 12 + k];
			rb.camera_matrix = cam;
			CUDA_CHECK_THROW(cudaDeviceSynchronize());
			const auto t0 = std::chrono::steady_clock::now();
			rb.render_buffer.resize({W, H});
			rb.render_buffer.reset_accumulation();
			for (int i = 0; i < specs[s].spp; ++i) tb.render_frame(rb.camera_matrix, rb.render_buffer, specs[s].srgb != 0);
			CUDA_CHECK_THROW(cudaDeviceSynchronize());
			if (s > 0 || p > 0) { seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); *n_timed_rays += (double)W * H * specs[s].spp; }   // first frame warms up
			CUDA_CHECK_THROW(cudaMemcpy2DFromArray(image.data(), (size_t)W * 4 * sizeof(float), rb.render_buffer.surface_provider().array(), 0, 0,
			                                       (size_t)W * 4 * sizeof(float), H, cudaMemcpyDeviceToHost));
			write_host(dir + "/" + prefix + "_" + std::to_string(s) + "_" + std::to_string(p) + ".f32", image.data(), image.size() * 4);
		}
	}
	return seconds;
}

// train, write the reference's own snapshots (internal/testbed.cu:127-141), render
static int run_snapshot(const std::string& dir, uint32_t calls, int W, int H, float fov_value, const std::string& specs, uint32_t batch) {
	Testbed tb;
	tb.load_model_config(dir + "/config.json");
	load_dataset(tb, dir);
	auto& b = tb.m_training_buffer;
	for (uint32_t i = 0; i < calls; ++i) tb.train(16, batch);
	CUDA_CHECK_THROW(cudaDeviceSynchronize());
	tb.save_snapshot(dir + "/ref_snapshot.msgpack", false);
	tb.save_snapshot(dir + "/ref_snapshot_opt.msgpack", true);
	double rays = 0;
	const double sec = render_specs(tb, dir, "ref_render", W, H, fov_value, parse_specs(specs), &rays);
	printf("{\"impl\": \"reference-cuda-sm100\", \"train_steps\": %u, \"loss\": %.6g, \"n_rays_per_batch\": %u, \"render_mrays_per_s\": %.3f}\n", calls * 16, (double)b.loss_scalar,
	       b.n_rays_per_batch, sec > 0 ? rays / sec / 1e6 : 0.0);
	return 0;
}

// Testbed::load_snapshot (internal/testbed.cu:143-163) of any msgpack file, then render; optionally continue training from it
static int run_loadrender(const std::string& dir, const std::string& file, int W, int H, float fov_value, const std::string& specs, uint32_t train_calls, uint32_t batch) {
	Testbed tb;
	if (train_calls) {
		// run.py's order: model config, data, then the snapshot (load_snapshot rebuilds the modules and puts the saved grid in place)
		tb.load_model_config(dir + "/config.json");
		load_dataset(tb, dir);
	}
	tb.load_snapshot(file);
	CUDA_CHECK_THROW(cudaDeviceSynchronize());
	double rays = 0;
	// specs "none": no rendering.  Needed for files WITH optimizer state: Ema::deserialize (ema.h:180) move-assigns a freshly
	// allocated buffer to m_weights_ema, the old one is freed, and Trainer::m_params_inference / the network's inference pointer
	// (set once in Trainer::initialize_params, trainer.h:91-105) keep pointing at it -- the reference cannot render after loading
	// such a file (use after free); training (use_inference_matrices = false) is unaffected.
	const double sec = specs == "none" ? 0.0 : render_specs(tb, dir, "ref_loadrender", W, H, fov_value, parse_specs(specs), &rays);
	double loss = -1, first_loss = -1;
	if (train_calls) {
		for (uint32_t i = 0; i < train_calls; ++i) { tb.train(16, batch); if (i == 0) first_loss = tb.m_training_buffer.loss_scalar; }
		loss = tb.m_training_buffer.loss_scalar;
		const size_t n = tb.m_network->n_params();
		write_device(dir + "/ref_loadtrain_params_f32.bin", tb.m_trainer->params(), n * 4);
	}
	printf("{\"impl\": \"reference-cuda-sm100\", \"render_mrays_per_s\": %.3f, \"render_ms_per_frame\": %.3f, \"timed_rays\": %.0f, \"trained_calls\": %u, \"first_call_loss\": %.6g, \"loss\": %.6g}\n",
	       sec > 0 ? rays / sec / 1e6 : 0.0, rays > 0 ? sec * 1e3 / (rays / ((double)W * H)) : 0.0, rays, train_calls, first_loss, loss);
	return 0;
}

// kernels of internal/testbed.cu (external linkage there; -rdc=true links them): launched directly so render_frame can be replayed stage by stage
__global__ void compact_rays_kernel(const uint32_t n_elements, RayPayload* src_payloads, Eigen::Array4f* src_rgba, RayPayload* dst_payloads, Eigen::Array4f* dst_rgba,
                                    RayPayload* final_payloads, Eigen::Array4f* final_rgba, uint32_t* n_counter, uint32_t* n_final_counter);
__global__ void shade_kernel(const uint32_t n_rays, const Eigen::Array4f* __restrict__ rgba, const RayPayload* __restrict__ payloads, Eigen::Array4f* __restrict__ frame_buffer,
                             const ERenderMode render_mode, const bool train_in_linear_colors);

// Testbed::render_frame (internal/testbed.cu:479-605) issued statement by statement through the reference's own module methods and kernels, with the buffers
// written out between the stages (the first `max_rounds` march rounds are dumped, all rounds are run)
static int run_renderdump(const std::string& dir, const std::string& file, int W, int H, float fov_value, int mode, uint32_t max_rounds) {
	using namespace tcnn;
	Testbed tb;
	tb.load_snapshot(file);
	tb.set_fov(fov_value);
	auto poses = read_file<float>(dir + "/test_xforms.f32");
	auto& rb = tb.m_rendering_buffer;
	Eigen::Matrix<float, 3, 4> camera_matrix;
	for (int k = 0; k < 12; ++k) camera_matrix.data()[k] = poses[k];
	rb.render_mode = (ERenderMode)mode;
	CudaRenderBuffer& render_buffer = rb.render_buffer;
	render_buffer.resize({W, H});
	render_buffer.reset_accumulation();
	cudaStream_t stream = tb.m_stream;

	const Vector2i resolution = render_buffer.resolution();
	const Vector2f focal_length = resolution[rb.fov_axis] * rb.relative_focal_length;
	const uint32_t n_pixels = resolution.x() * resolution.y();
	const ERenderMode render_mode = rb.render_mode;
	const uint32_t w = tb.m_network->padded_output_width();
	{
		const uint32_t n_elements = next_multiple(n_pixels, BATCH_SIZE_MULTIPLE);
		rb.rays[0].enlarge(n_elements); rb.rays[1].enlarge(n_elements); rb.rays_hit.enlarge(n_elements);
		rb.coords.enlarge(n_elements * rb.MAX_STEPS_INBETWEEN_COMPACTION);
		rb.rgbsigma.enlarge(n_elements * rb.MAX_STEPS_INBETWEEN_COMPACTION * w);
	}
	render_buffer.clear_frame_buffer(stream);
	tb.m_sampler->generate_rays_from_camera_matrix(resolution, focal_length, rb.principal_point, camera_matrix, tb.m_envmap.envmap->params_inference(), tb.m_envmap.resolution,
	                                               rb.rays[0].payload, render_buffer, stream);
	CUDA_CHECK_THROW(cudaMemsetAsync(rb.rays[0].rgba.data(), 0, rb.rays[0].rgba.get_bytes(), stream));
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	write_device(dir + "/ref_rd_bitfield_u8.bin", tb.m_sampler->density_grid_bitfield.data(), tb.m_sampler->density_grid_bitfield.size());
	write_device(dir + "/ref_rd_grid_f32.bin", tb.m_sampler->density_grid.data(), (size_t)tb.m_sampler->n_total_elements * 4);
	write_device(dir + "/ref_rd_payload_init.bin", rb.rays[0].payload.data(), (size_t)n_pixels * sizeof(RayPayload));
	CUDA_CHECK_THROW(cudaMemsetAsync(rb.n_hit_counter.data(), 0, sizeof(uint32_t), stream));
	uint32_t i_steps_offset = 1, double_buffer_index = 0, n_alive = n_pixels, round = 0;
	std::ofstream meta(dir + "/ref_rd_meta.txt");
	meta << "n_pixels " << n_pixels << "\npayload_bytes " << sizeof(RayPayload) << "\n";
	while (i_steps_offset < MARCH_ITER) {
		RaysSoA& rays_previous = rb.rays[double_buffer_index % 2];
		RaysSoA& rays_current = rb.rays[(double_buffer_index + 1) % 2];
		++double_buffer_index;
		CUDA_CHECK_THROW(cudaMemsetAsync(rb.n_alive_counter.data(), 0, sizeof(uint32_t), stream));
		linear_kernel(compact_rays_kernel, 0, stream, n_alive, rays_previous.payload.data(), rays_previous.rgba.data(), rays_current.payload.data(), rays_current.rgba.data(),
		              rb.rays_hit.payload.data(), rb.rays_hit.rgba.data(), rb.n_alive_counter.data(), rb.n_hit_counter.data());
		CUDA_CHECK_THROW(cudaMemcpyAsync(&n_alive, rb.n_alive_counter.data(), sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		if (n_alive == 0) break;
		const uint32_t n_steps = tcnn::clamp(n_pixels / n_alive, rb.MIN_STEPS_INBETWEEN_COMPACTION, rb.MAX_STEPS_INBETWEEN_COMPACTION);
		const bool dump = round < max_rounds;
		const std::string tag = dir + "/ref_rd_r" + std::to_string(round);
		if (dump) {
			write_device(tag + "_payload_in.bin", rays_current.payload.data(), (size_t)n_alive * sizeof(RayPayload));
			write_device(tag + "_rgba_in.bin", rays_current.rgba.data(), (size_t)n_alive * 16);
		}
		tb.m_sampler->generate_coords_from_rays(n_alive, n_steps, rays_current.payload, rb.coords, stream);
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		if (dump) {
			write_device(tag + "_payload_gen.bin", rays_current.payload.data(), (size_t)n_alive * sizeof(RayPayload));
			write_device(tag + "_coords.bin", rb.coords.data(), (size_t)n_alive * n_steps * sizeof(Coordinate));
		}
		const uint32_t n_elements = next_multiple(n_alive * n_steps, BATCH_SIZE_MULTIPLE);
		GPUMatrix<float> coords_matrix((float*)rb.coords.data(), sizeof(Coordinate) / sizeof(float), n_elements);
		GPUMatrix<precision_t> rgbsigma_matrix(rb.rgbsigma.data(), w, n_elements);
		tb.m_network->inference_mixed_precision(stream, coords_matrix, rgbsigma_matrix);
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		if (dump) write_device(tag + "_rgbsigma.bin", rb.rgbsigma.data(), (size_t)n_alive * n_steps * w * 2);
		if (render_mode == ERenderMode::Normals) {
			tb.m_network->input_gradient(stream, 3, coords_matrix, coords_matrix);
			CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
			if (dump) write_device(tag + "_coords_grad.bin", rb.coords.data(), (size_t)n_alive * n_steps * sizeof(Coordinate));
		}
		tb.m_render->volume_rendering(n_alive, n_steps, i_steps_offset, rays_current, rb.coords, rb.rgbsigma, w, tb.m_density_activation, tb.m_rgb_activation, render_mode,
		                              tb.m_sampler->aabb, tb.m_sampler->n_cascades, tb.m_sampler->min_cone_stepsize, camera_matrix, 1.f / tb.m_scene_scale, stream);
		CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
		if (dump) {
			write_device(tag + "_payload_out.bin", rays_current.payload.data(), (size_t)n_alive * sizeof(RayPayload));
			write_device(tag + "_rgba_out.bin", rays_current.rgba.data(), (size_t)n_alive * 16);
			meta << "round " << round << " " << n_alive << " " << n_steps << " " << i_steps_offset << "\n";
		}
		i_steps_offset += n_steps;
		++round;
	}
	uint32_t n_hit = 0;
	CUDA_CHECK_THROW(cudaMemcpyAsync(&n_hit, rb.n_hit_counter.data(), sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	write_device(dir + "/ref_rd_hit_payload.bin", rb.rays_hit.payload.data(), (size_t)n_hit * sizeof(RayPayload));
	write_device(dir + "/ref_rd_hit_rgba.bin", rb.rays_hit.rgba.data(), (size_t)n_hit * 16);
	linear_kernel(shade_kernel, 0, stream, n_hit, rb.rays_hit.rgba.data(), rb.rays_hit.payload.data(), render_buffer.frame_buffer(), render_mode, tb.m_render->train_in_linear_color);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	write_device(dir + "/ref_rd_frame.bin", render_buffer.frame_buffer(), (size_t)n_pixels * 16);
	render_buffer.set_color_space(tb.m_color_space);
	render_buffer.set_tonemap_curve(rb.tonemap_curve);
	render_buffer.accumulate(stream);
	render_buffer.tonemap(rb.exposure, tb.m_background_color, EColorSpace::SRGB, stream);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	write_device(dir + "/ref_rd_accum.bin", render_buffer.accumulate_buffer(), (size_t)n_pixels * 16);
	std::vector<float> image((size_t)W * H * 4);
	CUDA_CHECK_THROW(cudaMemcpy2DFromArray(image.data(), (size_t)W * 4 * sizeof(float), render_buffer.surface_provider().array(), 0, 0, (size_t)W * 4 * sizeof(float), H, cudaMemcpyDeviceToHost));
	write_host(dir + "/ref_rd_surface.f32", image.data(), image.size() * 4);
	meta << "rounds_total " << round << "\nn_hit " << n_hit << "\nn_cascades " << tb.m_sampler->n_cascades << "\nscene_scale " << tb.m_scene_scale << "\n";
	std::cout << "renderdump: " << n_pixels << " pixels, " << round << " rounds, " << n_hit << " hits" << std::endl;
	return 0;
}

// Testbed::load_training_data (internal/testbed.cu:95-125): NeRFSynthetic::load_from_json (stb_image, from_rgba32, spec_opengl_to_opencv) + sampler init
static int run_loaddata(const std::string& dir, const std::string& json_path, float scale) {
	Testbed tb;
	tb.load_model_config(dir + "/config.json");
	tb.load_training_data(json_path, scale, Vector3f::Constant(0.5f));
	CUDA_CHECK_THROW(cudaDeviceSynchronize());
	NeRFSynthetic& d = tb.m_training_buffer.dataset;
	write_device(dir + "/ref_ld_xforms.f32", d.xforms.data(), d.n_images * 12 * sizeof(float));
	write_device(dir + "/ref_ld_images.f16", d.images.data(), d.images.size() * 2);
	write_device(dir + "/ref_ld_grid0_f32.bin", tb.m_sampler->density_grid.data(), (size_t)tb.m_sampler->n_total_elements * 4);
	std::ofstream meta(dir + "/ref_ld_meta.txt");
	meta.precision(9);
	meta << d.n_images << " " << d.image_resolution.x() << " " << d.image_resolution.y() << " " << d.focal_length.x() << " " << d.focal_length.y() << " "
	     << d.principal_point.x() << " " << d.principal_point.y() << "\n";
	for (auto& f : d.filenames) meta << f << "\n";
	std::cout << "loaddata: " << d.n_images << " images " << d.image_resolution.x() << "x" << d.image_resolution.y() << std::endl;
	return 0;
}

// BASELINE configs[0]: the reference's GridEncoding alone, exactly as NGPNetwork builds it (ngp_network.h:162, 620-632)
static int run_c1(const std::string& dir, uint32_t iters, uint32_t log2_T) {
	using T = precision_t;
	auto xyz = read_file<float>(dir + "/c1_xyz.f32");
	auto table = read_file<uint16_t>(dir + "/c1_table.f16");
	auto dy = read_file<uint16_t>(dir + "/c1_dy.f16");
	const uint32_t n = (uint32_t)(xyz.size() / 3);
	const uint32_t n_levels = 16, base = 16;
	nlohmann::json cfg = {{"otype", "HashGrid"}, {"n_levels", n_levels}, {"n_features_per_level", 2}, {"log2_hashmap_size", log2_T}, {"base_resolution", base}};
	cfg["per_level_scale"] = std::exp(std::log(2048.0f / (float)base) / (n_levels - 1));
	std::unique_ptr<tcnn::Encoding<T>> enc(tcnn::create_encoding<T>(3, cfg, 16));
	const size_t P = enc->n_params();
	if (P != table.size() || dy.size() != (size_t)n * 32) throw std::runtime_error("c1: table / dy size mismatch");
	tcnn::GPUMemory<float> d_xyz(xyz.size()), d_full(P);
	tcnn::GPUMemory<T> d_params(P), d_grads(P), d_out((size_t)n * 32), d_dy((size_t)n * 32);
	tcnn::pcg32 rnd(1337);
	enc->initialize_params(rnd, d_full.data(), d_params.data(), d_params.data(), d_params.data(), d_grads.data());
	CUDA_CHECK_THROW(cudaMemcpy(d_params.data(), table.data(), P * sizeof(T), cudaMemcpyHostToDevice));
	CUDA_CHECK_THROW(cudaMemcpy(d_xyz.data(), xyz.data(), xyz.size() * sizeof(float), cudaMemcpyHostToDevice));
	CUDA_CHECK_THROW(cudaMemcpy(d_dy.data(), dy.data(), dy.size() * sizeof(T), cudaMemcpyHostToDevice));
	cudaStream_t stream; CUDA_CHECK_THROW(cudaStreamCreate(&stream));
	cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
	auto fwd = [&]() { enc->encode(stream, n, {d_xyz.data(), 3}, {d_out.data(), 32}, nullptr, false); };
	auto bwd = [&]() { enc->backward(stream, n, {d_dy.data(), 32}, nullptr, {}, {d_xyz.data(), 3}, false); };
	auto timed = [&](auto&& fn) {
		for (int i = 0; i < 10; ++i) fn();
		std::vector<float> ms(iters);
		for (uint32_t i = 0; i < iters; ++i) { cudaEventRecord(a, stream); fn(); cudaEventRecord(b, stream); cudaEventSynchronize(b); cudaEventElapsedTime(&ms[i], a, b); }
		std::sort(ms.begin(), ms.end());
		return ms[iters / 2] * 1e3f;
	};
	const float fwd_us = timed(fwd), bwd_us = timed(bwd);
	CUDA_CHECK_THROW(cudaStreamSynchronize(stream));
	write_device(dir + "/ref_c1_enc.f16", d_out.data(), (size_t)n * 32 * sizeof(T));
	write_device(dir + "/ref_c1_grad.f16", d_grads.data(), P * sizeof(T));
	printf("{\"impl\": \"reference-cuda-sm100 GridEncoding\", \"n_points\": %u, \"n_params\": %zu, \"iterations\": %u, \"fwd_us\": %.3f, \"bwd_us\": %.3f}\n", n, P, iters, fwd_us, bwd_us);
	return 0;
}

int main(int argc, char** argv) {
	try {
		if (argc >= 3 && std::string(argv[1]) == "dump") return run_dump(argv[2], argc > 3 ? (uint32_t)atoi(argv[3]) : (1u << 18), argc > 4 ? (uint32_t)atoi(argv[4]) : 0u);
		if (argc >= 4 && std::string(argv[1]) == "c1") return run_c1(argv[2], (uint32_t)atoi(argv[3]), argc > 4 ? (uint32_t)atoi(argv[4]) : 19u);
		if (argc >= 5 && std::string(argv[1]) == "bench") return run_bench(argv[2], (uint32_t)atoi(argv[3]), (uint32_t)atoi(argv[4]), argc > 5 ? (uint32_t)atoi(argv[5]) : (1u << 18));
		const std::string cmd = argc > 1 ? argv[1] : "";
		if (argc >= 5 && cmd == "benchsteps")
			return run_benchsteps(argv[2], (uint32_t)atoi(argv[3]), (uint32_t)atoi(argv[4]), argc > 5 ? (uint32_t)atoi(argv[5]) : (1u << 18), argc > 6 ? atoi(argv[6]) : 0,
			                      argc > 7 ? atoi(argv[7]) : 0, argc > 8 ? atoi(argv[8]) : 0, argc > 9 ? (float)atof(argv[9]) : 0.6911112f);
		if (argc >= 8 && cmd == "snapshot") return run_snapshot(argv[2], (uint32_t)atoi(argv[3]), atoi(argv[4]), atoi(argv[5]), (float)atof(argv[6]), argv[7], argc > 8 ? (uint32_t)atoi(argv[8]) : (1u << 18));
		if (argc >= 8 && cmd == "loadrender")
			return run_loadrender(argv[2], argv[3], atoi(argv[4]), atoi(argv[5]), (float)atof(argv[6]), argv[7], argc > 8 ? (uint32_t)atoi(argv[8]) : 0u, argc > 9 ? (uint32_t)atoi(argv[9]) : (1u << 18));
		if (argc >= 9 && cmd == "renderdump") return run_renderdump(argv[2], argv[3], atoi(argv[4]), atoi(argv[5]), (float)atof(argv[6]), atoi(argv[7]), (uint32_t)atoi(argv[8]));
		if (argc >= 5 && cmd == "loaddata") return run_loaddata(argv[2], argv[3], (float)atof(argv[4]));
		if (argc >= 8 && std::string(argv[1]) == "trainrender")
			return run_trainrender(argv[2], (uint32_t)atoi(argv[3]), atoi(argv[4]), atoi(argv[5]), atoi(argv[6]), (float)atof(argv[7]), argc > 8 ? (uint32_t)atoi(argv[8]) : (1u << 18));
		std::cerr << "usage: ref_harness dump <dir> [batch] [n_rays] | bench <dir> <calls> <warmup_calls> [batch] | trainrender <dir> <calls> <W> <H> <spp> <fov> [batch]" << std::endl;
		return 2;
	} catch (const std::exception& e) {
		std::cerr << "ref_harness: " << e.what() << std::endl;
		return 1;
	}
}
