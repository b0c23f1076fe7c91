// pybind11 module `pyngp` -- same classes, enums, method names, argument names and defaults as the reference's
// internal/python_api.cu:48-138, so experiment/nerf_synthetic/python/run.py runs against it unchanged.  Eigen casters are
// replaced by numpy / sequence conversions that accept the same Python values ([0.5,0.5,0.5], a 3x4 or 4x4 np.matrix).
#include <pybind11/pybind11.h>
#include <pybind11/numpy.h>
#include <pybind11/stl.h>
#include "testbed.h"

namespace py = pybind11;
using namespace ngp;

static Mat34 to_mat34(const py::object& o) {
	py::array_t<float, py::array::c_style | py::array::forcecast> a = py::array_t<float, py::array::c_style | py::array::forcecast>::ensure(o);
	if (!a || a.ndim() != 2 || a.shape(0) < 3 || a.shape(1) != 4) throw std::runtime_error("expected a 3x4 (or 4x4) camera matrix");
	Mat34 m;
	auto r = a.unchecked<2>();
	for (int i = 0; i < 3; ++i) for (int j = 0; j < 4; ++j) m.at(i, j) = r(i, j);
	return m;
}

static void to_vec(const py::object& o, float* dst, size_t n, const char* what) {
	py::sequence s = py::reinterpret_borrow<py::sequence>(py::module_::import("numpy").attr("asarray")(o).attr("reshape")(-1).attr("tolist")());
	if (s.size() != n) throw std::runtime_error(std::string(what) + ": wrong number of elements");
	for (size_t i = 0; i < n; ++i) dst[i] = s[i].cast<float>();
}

PYBIND11_MODULE(pyngp, m) {
	m.def("shard_range", [](size_t n_params, uint32_t rank, uint32_t world) {
		if (world < 1 || world > NGP_MAX_PEERS || rank >= world) throw std::runtime_error("shard_range: bad rank / world");
		const ShardRange r = shard_range(n_params, rank, world);
		return py::make_tuple(r.begin, r.end, r.size);
	}, "(begin, end, shard size) of the parameter range a rank owns in the sharded data-parallel optimizer");
	m.def("decode_png", [](const std::string& path) {
		int w = 0, h = 0;
		const std::vector<uint8_t> px = decode_png_file(path, w, h);
		py::array_t<uint8_t> img({h, w, 4});
		std::memcpy(img.request().ptr, px.data(), px.size());
		return img;
	}, "The dataset loader's PNG decoder (host code): RGBA8 [h][w][4], converted as stb_image's stbi_load(path, ..., 4) converts.");
	m.doc() = "Instant neural graphics primitives (NeRF) minimal version -- B200-native build";

	py::enum_<EActivation>(m, "Activation")
		.value("None", EActivation::None).value("ReLU", EActivation::ReLU).value("Logistic", EActivation::Logistic).value("Exponential", EActivation::Exponential)
		.export_values();
	py::enum_<ERenderMode>(m, "RenderMode")
		.value("AO", ERenderMode::AO).value("Shade", ERenderMode::Shade).value("Normals", ERenderMode::Normals).value("Depth", ERenderMode::Depth)
		.value("Distance", ERenderMode::Distance).value("Stepsize", ERenderMode::Stepsize).value("Cost", ERenderMode::Cost)
		.export_values();
	py::enum_<EColorSpace>(m, "ColorSpace").value("Linear", EColorSpace::Linear).value("SRGB", EColorSpace::SRGB).export_values();
	py::enum_<ETonemapCurve>(m, "TonemapCurve")
		.value("Identity", ETonemapCurve::Identity).value("ACES", ETonemapCurve::ACES).value("Hable", ETonemapCurve::Hable).value("Reinhard", ETonemapCurve::Reinhard)
		.export_values();
	py::enum_<ELossType>(m, "LossType")
		.value("L2", ELossType::L2).value("L1", ELossType::L1).value("Mape", ELossType::Mape).value("Smape", ELossType::Smape)
		.value("SmoothL1", ELossType::SmoothL1).value("LogL1", ELossType::LogL1).value("RelativeL2", ELossType::RelativeL2)
		.export_values();

	py::class_<Testbed> testbed(m, "Testbed");
	testbed
		.def(py::init())
		.def("load_model_config", &Testbed::load_model_config, "Reload the network from a config file.")
		.def("save_snapshot", &Testbed::save_snapshot, py::arg("snapshot_path"), py::arg("serialize_optimizer") = false, "Save a snapshot of the currently trained model")
		.def("load_snapshot", &Testbed::load_snapshot, py::arg("snapshot_path"), "Load a previously saved snapshot")
		.def("load_training_data", [](Testbed& t, const std::string& path, float scale, const py::object& offset) {
			float off[3]; to_vec(offset, off, 3, "offset");
			t.load_training_data(path, scale, off);
		}, "Load training data from a given path.")
		.def("load_training_data_from_memory", [](Testbed& t, py::array_t<uint8_t, py::array::c_style> rgba8, py::array_t<float, py::array::c_style | py::array::forcecast> xforms,
		                                          float camera_angle_x, float scale, const py::object& offset) {
			if (rgba8.ndim() != 4 || rgba8.shape(3) != 4) throw std::runtime_error("rgba8 must be [n][H][W][4] uint8");
			if (xforms.ndim() != 3 || xforms.shape(0) != rgba8.shape(0) || xforms.shape(1) < 3 || xforms.shape(2) != 4) throw std::runtime_error("xforms must be [n][3|4][4]");
			float off[3]; to_vec(offset, off, 3, "offset");
			const uint32_t n = (uint32_t)rgba8.shape(0);
			std::vector<Mat34> m(n);
			auto r = xforms.unchecked<3>();
			for (uint32_t i = 0; i < n; ++i) for (int a = 0; a < 3; ++a) for (int b = 0; b < 4; ++b) m[i].at(a, b) = r(i, a, b);
			const uint8_t* ptr = rgba8.data();
			py::gil_scoped_release rel;
			t.load_training_data_from_memory(n, (int)rgba8.shape(2), (int)rgba8.shape(1), camera_angle_x, m.data(), ptr, scale, off);
		}, py::arg("rgba8"), py::arg("xforms"), py::arg("camera_angle_x"), py::arg("scale") = 0.33f, py::arg("offset") = std::vector<float>{0.5f, 0.5f, 0.5f},
		   "Same as load_training_data for images already decoded in host memory.")
		.def("load_procedural_training_data", [](Testbed& t, uint32_t n_images, int width, int height, uint32_t seed, float scale, const py::object& offset) {
			float off[3]; to_vec(offset, off, 3, "offset");
			t.load_procedural_training_data(n_images, width, height, seed, scale, off);
		}, py::arg("n_images") = 100, py::arg("width") = 800, py::arg("height") = 800, py::arg("seed") = 2024, py::arg("scale") = 0.33f,
		   py::arg("offset") = std::vector<float>{0.5f, 0.5f, 0.5f}, "Benchmark fixture: procedural nerf_synthetic-shaped scene generated on the device.")
		.def("train", [](Testbed& t, uint32_t n_steps, uint32_t batch) { py::gil_scoped_release rel; t.train(n_steps, batch); }, "Perform a specified number of training steps.")
		.def("render", [](Testbed& t, int width, int height, int spp, bool to_srgb) {
			py::array_t<float> image({height, width, 4});
			float* ptr = (float*)image.request().ptr;
			{ py::gil_scoped_release rel; t.render_to_host(width, height, spp, to_srgb, ptr); }
			return image;
		}, "Renders an image at the requested resolution.", py::arg("width") = 1920, py::arg("height") = 1080, py::arg("spp") = 1, py::arg("to_srgb") = true)
		.def("render_band", [](Testbed& t, int width, int height, int spp, bool to_srgb) {
			const int r0 = t.render_row_begin(height), r1 = t.render_row_end(height);
			py::array_t<float> image({r1 - r0, width, 4});
			float* ptr = (float*)image.request().ptr;
			{ py::gil_scoped_release rel; t.render_band_to_host(width, height, spp, to_srgb, ptr); }
			return image;
		}, "Renders the rows set by set_render_rows and returns them alone, [rows][width][4].", py::arg("width") = 1920, py::arg("height") = 1080, py::arg("spp") = 1, py::arg("to_srgb") = true)
		.def("render_frames_ms", [](Testbed& t, int width, int height, int spp, bool to_srgb) { py::gil_scoped_release rel; return t.render_frames_ms(width, height, spp, to_srgb); },
		     "Measurement hook: renders spp frames into the device-side surface (no host copy) and returns the wall-clock milliseconds.",
		     py::arg("width") = 1920, py::arg("height") = 1080, py::arg("spp") = 1, py::arg("to_srgb") = true)
		.def("envmap_image", [](const Testbed& t) {
			// the trained environment map as render() uses it (EMA weights), [resolution[1]][resolution[0]][4] floats; zeros when not trained
			const auto& e = t.m_envmap;
			py::array_t<float> img({e.resolution[1], e.resolution[0], 4});
			float* out = (float*)img.request().ptr;
			const size_t n = (size_t)e.resolution[0] * e.resolution[1] * 4;
			if (!t.m_train_envmap || e.n_params != n) { std::fill(out, out + n, 0.f); return img; }
			std::vector<uint16_t> h(n);
			NGP_CUDA_CHECK(cudaMemcpy(h.data(), e.params_inference(), n * 2, cudaMemcpyDeviceToHost));
			for (size_t k = 0; k < n; ++k) {
				const uint32_t s = (h[k] >> 15) & 1, ex = (h[k] >> 10) & 0x1f, m = h[k] & 0x3ff;
				const float v = ex == 0 ? std::ldexp((float)m, -24) : ex == 31 ? (m ? NAN : INFINITY) : std::ldexp((float)(m | 0x400), (int)ex - 25);
				out[k] = s ? -v : v;
			}
			return img;
		}, "The trained environment map (inference weights) as float32 [y][x][4].")
		.def("set_render_rows", &Testbed::set_render_rows,
		     "Tile-sharded inference: render() marches only the rays of rows [row_begin, row_end) (the other rows of the returned image hold the background); "
		     "row_end = -1 restores the whole image.  Bands are bit-identical to the same rows of the full frame.", py::arg("row_begin"), py::arg("row_end"))
		.def_readwrite("loss_type", &Testbed::m_loss_type)
		.def_readwrite("density_activation", &Testbed::m_density_activation)
		.def_readwrite("rgb_activation", &Testbed::m_rgb_activation)
		.def_property("background_color",
			[](const Testbed& t) { return std::vector<float>(t.m_background_color, t.m_background_color + 4); },
			[](Testbed& t, const py::object& v) { to_vec(v, t.m_background_color, 4, "background_color"); })
		.def_readwrite("color_space", &Testbed::m_color_space)
		.def_readonly("training_buffer", &Testbed::m_training_buffer)
		.def_readonly("rendering_buffer", &Testbed::m_rendering_buffer)
		.def_readwrite("shard_grid_refresh", &Testbed::m_shard_grid_refresh, "data parallel: shard the occupancy-grid refresh over the ranks (default on; off = every rank refreshes everything)")
		.def_property("coarse_skip", [](const Testbed& t) { return t.coarse_skip(); }, [](Testbed& t, bool v) { t.set_coarse_skip(v); },
		              "inference march crosses empty 8x8x8 blocks of the occupancy grid in one hop (default on; results are identical either way)")
		.def_readwrite("pipeline_march", &Testbed::m_pipeline_march, "march step i+1 on a second stream underneath step i (default on)")
		.def_property("fov", &Testbed::fov, &Testbed::set_fov)
		.def_property("fov_xy",
			[](const Testbed& t) { return std::vector<float>{focal_length_to_fov(1, t.m_rendering_buffer.relative_focal_length[0]), focal_length_to_fov(1, t.m_rendering_buffer.relative_focal_length[1])}; },
			[](Testbed& t, const py::object& v) { float f[2]; to_vec(v, f, 2, "fov_xy"); for (int k = 0; k < 2; ++k) t.m_rendering_buffer.relative_focal_length[k] = fov_to_focal_length(1, f[k]); })
		.def("set_nerf_camera_matrix", [](Testbed& t, const py::object& cam) { t.set_nerf_camera_matrix(to_mat34(cam)); })
		.def("set_camera_to_training_view", &Testbed::set_camera_to_training_view)
		.def("training_image", [](const Testbed& t, uint32_t i) {
			const NeRFSynthetic& d = t.m_training_buffer.dataset;
			py::array_t<float> image({d.image_resolution[1], d.image_resolution[0], 4});
			t.training_image_to_host(i, (float*)image.request().ptr);
			return image;
		})
		.def_property_readonly("training_xforms", [](const Testbed& t) {
			// camera-to-world matrices after spec_opengl_to_opencv, [n][12] column-major (what nerf_synthetic.cu:98 uploads)
			const NeRFSynthetic& d = t.m_training_buffer.dataset;
			py::array_t<float> out({(py::ssize_t)d.n_images, (py::ssize_t)12});
			if (d.n_images) NGP_CUDA_CHECK(cudaMemcpy(out.request().ptr, d.xforms.data(), (size_t)d.n_images * sizeof(Mat34), cudaMemcpyDeviceToHost));
			return out;
		})
		.def_property_readonly("training_focal_length", [](const Testbed& t) { const NeRFSynthetic& d = t.m_training_buffer.dataset; return std::vector<float>{d.focal_length[0], d.focal_length[1]}; })
		.def_property_readonly("training_filenames", [](const Testbed& t) { return t.m_training_buffer.dataset.filenames; })
		.def_property_readonly("n_training_images", [](const Testbed& t) { return t.m_training_buffer.dataset.n_images; })
		// ---- additions for data-parallel training (not in the reference) ----
		.def("set_data_parallel", [](Testbed& t, int rank, int world, uintptr_t comm, bool peer_memory) { t.set_data_parallel(rank, world, (void*)comm, peer_memory); },
		     py::arg("rank"), py::arg("world_size"), py::arg("nccl_comm"), py::arg("peer_memory") = false,
		     "Data-parallel training, one process per GPU (collective).  peer_memory: fused reduce-scatter + Adam + all-gather over NVLink peer memory.")
		.def_property_readonly("peer_memory_enabled", &Testbed::peer_memory_enabled)
		.def("gather_optimizer_state", &Testbed::gather_optimizer_state, "all-gather sharded optimizer state (collective); call before save_snapshot(..., True) in peer-memory mode")
		.def_property_readonly("gradients_ptr", [](const Testbed& t) { return (uintptr_t)t.gradients(); })
		.def_property_readonly("n_params", &Testbed::n_params)
		.def_property_readonly("stream_ptr", [](const Testbed& t) { return (uintptr_t)t.m_stream; })
		;

	py::class_<Testbed::TrainingBuffer>(testbed, "TrainingBuffer")
		.def_readonly("n_rays_per_batch", &Testbed::TrainingBuffer::n_rays_per_batch)
		.def_readonly("measured_batch_size_before_compaction", &Testbed::TrainingBuffer::measured_batch_size_before_compaction)
		.def_readonly("measured_batch_size", &Testbed::TrainingBuffer::measured_batch_size)
		.def_readonly("loss", &Testbed::TrainingBuffer::loss_scalar)
		.def_readonly("training_prep_ms", &Testbed::TrainingBuffer::training_prep_ms)
		.def_readonly("training_ms", &Testbed::TrainingBuffer::training_ms)
		.def_readonly("enqueue_ms", &Testbed::TrainingBuffer::enqueue_ms)
		.def_readonly("i_step", &Testbed::TrainingBuffer::i_step)
		// per-step device counters of the last train() call (exact, for throughput accounting)
		.def_readonly("n_coords_counter_cpu", &Testbed::TrainingBuffer::n_coords_counter_cpu)
		.def_readonly("n_compacted_coords_counter_cpu", &Testbed::TrainingBuffer::n_compacted_coords_counter_cpu)
		;

	py::class_<Testbed::RenderingBuffer>(testbed, "RenderingBuffer")
		.def_readwrite("render_mode", &Testbed::RenderingBuffer::render_mode)
		.def_readwrite("tonemap_curve", &Testbed::RenderingBuffer::tonemap_curve)
		.def_readwrite("exposure", &Testbed::RenderingBuffer::exposure)
		.def_readwrite("fov_axis", &Testbed::RenderingBuffer::fov_axis)
		.def_property("relative_focal_length",
			[](const Testbed::RenderingBuffer& r) { return std::vector<float>{r.relative_focal_length[0], r.relative_focal_length[1]}; },
			[](Testbed::RenderingBuffer& r, const py::object& v) { to_vec(v, r.relative_focal_length, 2, "relative_focal_length"); })
		.def_property("principal_point",
			[](const Testbed::RenderingBuffer& r) { return std::vector<float>{r.principal_point[0], r.principal_point[1]}; },
			[](Testbed::RenderingBuffer& r, const py::object& v) { to_vec(v, r.principal_point, 2, "principal_point"); })
		.def_readwrite("device_loop", &Testbed::RenderingBuffer::device_loop, "render rounds decided on the device, the live-ray count read back every ROUNDS_PER_READBACK rounds (default on; off: one host synchronisation per round like the reference)")
		.def_readwrite("CONCURRENT_BANDS", &Testbed::RenderingBuffer::CONCURRENT_BANDS, "device loop: 2 = march the frame as two row bands on two streams at once (default), 1 = one pass")
		.def_readwrite("CONCURRENT_BANDS_MIN_PIXELS", &Testbed::RenderingBuffer::CONCURRENT_BANDS_MIN_PIXELS)
		.def_readwrite("MIN_BATCH_SAMPLES", &Testbed::RenderingBuffer::MIN_BATCH_SAMPLES, "device loop: a round's step count aims at max(pixels, this) samples (0 = the reference's rule)")
		.def_readwrite("fused_compaction", &Testbed::RenderingBuffer::fused_compaction, "device loop: compositing and the next round's compaction in one launch (default on)")
		.def_readwrite("ROUNDS_PER_READBACK", &Testbed::RenderingBuffer::ROUNDS_PER_READBACK)
		.def_readwrite("TAIL_STEPS", &Testbed::RenderingBuffer::TAIL_STEPS, "late rounds (batch below a quarter of the frame) take up to this many steps; <= MAX_STEPS_INBETWEEN_COMPACTION switches it off")
		.def_readonly("rounds_last_frame", &Testbed::RenderingBuffer::rounds_last_frame)
		.def_readonly("readbacks_last_frame", &Testbed::RenderingBuffer::readbacks_last_frame)
		.def_readwrite("MIN_STEPS_INBETWEEN_COMPACTION", &Testbed::RenderingBuffer::MIN_STEPS_INBETWEEN_COMPACTION)
		.def_readwrite("MAX_STEPS_INBETWEEN_COMPACTION", &Testbed::RenderingBuffer::MAX_STEPS_INBETWEEN_COMPACTION)
		;

	// NCCL bootstrap helpers for the data-parallel path: rank 0 makes the id, torch.distributed broadcasts the bytes
	m.def("nccl_unique_id", []() { return py::bytes(ngp_host_nccl_unique_id()); });
	m.def("nccl_comm_init", [](const py::bytes& id, int rank, int world) { return (uintptr_t)ngp_host_nccl_comm_init(std::string(id), rank, world); });
	m.def("nccl_comm_destroy", [](uintptr_t comm) { ngp_host_nccl_comm_destroy((void*)comm); });
}
