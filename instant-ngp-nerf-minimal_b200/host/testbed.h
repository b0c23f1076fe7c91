// Testbed -- same public surface as the reference's internal/testbed.h:30-204 (load_model_config, load_training_data,
// save/load_snapshot, train(n_steps, batch), render_frame / render, fov helpers, training_buffer / rendering_buffer views),
// driving the sampler / network / render modules whose kernels live behind the C ABI.
#pragma once
#include <algorithm>
#include <chrono>
#include <memory>
#include "common.h"
#include "json.h"
#include "dataset.h"
#include "occupancy_sampler.h"
#include "ngp_network.h"
#include "ray_marcher.h"
#include "render_buffer.h"

namespace ngp {

class Testbed {
public:
	Testbed();
	~Testbed();

	// ---- configuration & I/O (testbed.cu:44-163) ----
	void load_model_config(const std::string& config_path);
	void load_model_config_json(const Json& config);
	void save_snapshot(const std::string& snapshot_path, bool serialize_optimizer);
	void load_snapshot(const std::string& snapshot_path);
	void load_training_data(const std::string& data_path, float scale, const float offset[3]);
	// images already decoded in HOST memory (rgba8 sRGB straight alpha [n][H][W][4], OpenGL camera-to-world 3x4 per view): the
	// same upload + from_rgba32 conversion + sampler init as load_training_data, minus the file parsing
	void load_training_data_from_memory(uint32_t n_images, int width, int height, float camera_angle_x, const Mat34* xforms_ogl, const uint8_t* rgba8,
	                                    float scale, const float offset[3]);
	// benchmark fixture (no file I/O): procedural nerf_synthetic-shaped scene generated on the device
	void load_procedural_training_data(uint32_t n_images, int width, int height, uint32_t seed, float scale, const float offset[3]);

	// ---- training (testbed.cu:168-418) ----
	void train(uint32_t n_training_steps, uint32_t batch_size);
	void train_prep_nerf(uint32_t batch_size, uint32_t n_training_steps, cudaStream_t stream);
	void train_nerf(uint32_t target_batch_size, uint32_t n_training_steps, cudaStream_t stream);
	void train_nerf_step(uint32_t target_batch_size, uint32_t n_rays_per_batch, uint32_t* counters /* [2]: rays, samples */, uint32_t* compacted_counter,
	                     float* loss, cudaStream_t stream);

	// ---- rendering (testbed.cu:479-605, python_api.cu:22-46) ----
	void render_frame(const Mat34& camera_matrix, CudaRenderBuffer& render_buffer, bool to_srgb = true);
	// renders spp frames and copies the float4 image into `out` ([height][width][4] floats, host memory)
	void render_to_host(int width, int height, int spp, bool to_srgb, float* out);
	// tile-sharded inference: this rank renders rows [row_begin, row_end) only (no collective)
	void set_render_rows(int row_begin, int row_end) { m_render_row_begin = row_begin; m_render_row_end = row_end; }
	int render_row_begin(int height) const { return m_render_row_end < 0 ? 0 : std::max(0, std::min(m_render_row_begin, height)); }
	int render_row_end(int height) const { return m_render_row_end < 0 ? height : std::max(render_row_begin(height), std::min(m_render_row_end, height)); }
	void render_band_to_host(int width, int height, int spp, bool to_srgb, float* out_band);
	// measurement hook: spp frames rendered into the device-side surface (ray generation -> tonemap, what the metric's Mrays/s counts,
	// SURVEY.md 8d), no device -> host copy; returns the wall-clock milliseconds between two device synchronisations
	float render_frames_ms(int width, int height, int spp, bool to_srgb);

	float fov() const { return focal_length_to_fov(1, m_rendering_buffer.relative_focal_length[m_rendering_buffer.fov_axis]); }
	void set_fov(float val) { m_rendering_buffer.relative_focal_length[0] = m_rendering_buffer.relative_focal_length[1] = fov_to_focal_length(1, val); }
	// evaluation helpers: look through training camera i / fetch its ground-truth image (linear premultiplied RGBA, float)
	void set_camera_to_training_view(uint32_t i);
	void training_image_to_host(uint32_t i, float* out) const;
	void set_nerf_camera_matrix(const Mat34& cam) { m_rendering_buffer.camera_matrix = spec_opengl_to_opencv(cam, m_scene_scale, m_scene_offset); }

	// ---- data-parallel training (new: the reference is single-GPU).  Gradients are summed over ranks with NCCL
	//      before the optimizer step; each rank draws its own rays.  comm is an ncclComm_t created by the caller. ----
	// peer_memory: fused reduce-scatter + Adam + all-gather over NVLink peer memory (CUDA IPC between the ranks' processes)
	// instead of NCCL all-reduce + replicated optimizer.  Collective: every rank must call it.
	void set_data_parallel(int rank, int world_size, void* nccl_comm, bool peer_memory = false);
	bool peer_memory_enabled() const { return m_peer_enabled; }
	bool coarse_skip() const { return m_sampler ? m_sampler->coarse_skip : true; }
	void set_coarse_skip(bool v) { if (!m_sampler) throw std::runtime_error("coarse_skip: load_model_config first"); m_sampler->coarse_skip = v; }
	// all-gather of sharded optimizer state (collective; needed before a snapshot with serialize_optimizer in peer mode)
	void gather_optimizer_state();
	precision_t* gradients() const { return m_trainer->gradients(); }
	size_t n_params() const { return m_trainer->n_params(); }

public:
	uint32_t m_seed = 43;
	default_rng_t m_rng = default_rng_t{(uint64_t)43};
	cudaStream_t m_stream = nullptr;

	Json m_model_config;
	std::shared_ptr<OccupancySampler> m_sampler;
	std::shared_ptr<NGPNetwork> m_network;
	std::shared_ptr<RayMarcher> m_render;
	ELossType m_loss_type = ELossType::SmoothL1;
	std::shared_ptr<Trainer> m_trainer;
	EActivation m_density_activation = EActivation::Exponential;
	EActivation m_rgb_activation = EActivation::Logistic;
	float m_background_color[4] = {0.0f, 0.0f, 0.0f, 1.0f};
	EColorSpace m_color_space = EColorSpace::Linear;
	bool m_train_envmap = false;
	// trainable environment map (testbed.cu:82-90: TrainableBuffer<4, 2> + its own Ema(ExponentialDecay(Adam)) and loss).  The
	// reference stores the resolution as {height, width} and indexes x + y * resolution.x (utils/envmap.h:19-47): kept.
	struct Envmap {
		int resolution[2] = {256, 256};
		ELossType loss_type = ELossType::RelativeL2;
		OptimizerConfig opt;
		size_t n_params = 0;
		DeviceBuffer<float> params_fp32, m1, m2;
		DeviceBuffer<precision_t> params, gradients, ema;
		DeviceBuffer<uint32_t> param_steps;
		uint32_t step = 0; float lr_factor = 1.0f;
		precision_t* params_inference() const { return opt.has_ema ? ema.data() : params.data(); }
	} m_envmap;
	void envmap_optimizer_step(cudaStream_t stream, float loss_scale);      // Trainer::optimizer_step of the envmap (testbed.cu:296-298)

	float m_scene_scale = 1.f;
	float m_scene_offset[3] = {0.5f, 0.5f, 0.5f};

	struct TrainingBuffer {
		NeRFSynthetic dataset;
		uint32_t n_rays_per_batch = 1 << 12;
		uint32_t n_rays_total = 0;
		DeviceBuffer<Position> grid_positions;
		DeviceBuffer<uint32_t> grid_indices;
		// Rays and marched samples are double-buffered: the march of step i+1 only depends on the occupancy bitfield and the host
		// generator, so it runs on a second stream underneath step i's network / optimizer kernels.
		struct MarchSlot {
			DeviceBuffer<Ray> rays;
			DeviceBuffer<float> rays_rgba;
			DeviceBuffer<uint32_t> ray_coord_indices;
			DeviceBuffer<Coordinate> coords;
		} march[2];
		DeviceBuffer<Coordinate> coords_compacted;
		DeviceBuffer<uint32_t> compacted_src_index;   // marched-sample index of every compacted sample
		DeviceBuffer<precision_t> rgbsigma, dloss_drgbsigma;
		DeviceBuffer<uint32_t> n_coords_counter;            // [n_steps][2]: rays kept, samples counted
		DeviceBuffer<uint32_t> n_compacted_coords_counter;  // [n_steps]
		std::vector<uint32_t> n_coords_counter_cpu, n_compacted_coords_counter_cpu;
		uint32_t measured_batch_size_before_compaction = 0xffffffffu;
		uint32_t measured_batch_size = 0xffffffffu;
		DeviceBuffer<float> loss;
		DeviceBuffer<float> loss_sum;                       // reduce_sum result + scratch
		float loss_scalar = 0.f;
		float training_prep_ms = 0, training_ms = 0;
		float enqueue_ms = 0;                               // host time spent issuing the last train() call's work (before the one sync)
		uint32_t i_step = 0;
	} m_training_buffer;

	struct RenderingBuffer {
		ERenderMode render_mode = ERenderMode::Shade;
		ETonemapCurve tonemap_curve = ETonemapCurve::Identity;
		float exposure = 0.f;
		CudaRenderBuffer render_buffer;
		uint32_t fov_axis = 1;
		float relative_focal_length[2] = {1.f, 1.f};
		float principal_point[2] = {0.5f, 0.5f};
		Mat34 camera_matrix;
		RaysSoA rays[2];
		RaysSoA rays_hit;
		DeviceBuffer<uint32_t> n_hit_counter{1};
		DeviceBuffer<uint32_t> n_alive_counter{64};
		DeviceBuffer<Coordinate> coords;
		DeviceBuffer<precision_t> rgbsigma;
		// device-driven loop: the frame is marched as up to two row bands on two streams at once (one band's march and compositing
		// run under the other band's network kernels); every band owns its ray lists, sample buffers, counters and network scratch
		struct Band {
			RaysSoA rays[2]; RaysSoA rays_hit;
			DeviceBuffer<uint32_t> n_hit_counter{1}, round_ctl{NGP_RENDER_CTL_WORDS};
			DeviceBuffer<Coordinate> coords; DeviceBuffer<precision_t> rgbsigma; DeviceBuffer<uint8_t> workspace;
		} band[2];
		uint32_t CONCURRENT_BANDS = 2;         // ... when the frame (or band) has at least CONCURRENT_BANDS_MIN_PIXELS pixels: below that the rounds are latency bound and two streams only add launches
		uint32_t CONCURRENT_BANDS_MIN_PIXELS = 1u << 20;
		uint32_t MIN_STEPS_INBETWEEN_COMPACTION = 1;
		uint32_t MAX_STEPS_INBETWEEN_COMPACTION = 8;
		// Not in the reference: how the rounds are scheduled (the image does not depend on it).
		bool device_loop = true;              // rounds decided on the device, live-ray count read back every ROUNDS_PER_READBACK rounds (false: testbed.cu's loop, one sync per round)
		bool fused_compaction = true;         // device loop: the compositing launch also sorts its rays into the next round's buffer / the hit list
		uint32_t ROUNDS_PER_READBACK = 8;
		uint32_t MIN_BATCH_SAMPLES = 1u << 21; // the step count of a round aims at max(pixels, this) samples (0: the reference's rule, one sample per pixel)
		uint32_t TAIL_STEPS = 32;             // late rounds (batch below a quarter of the frame) take up to this many steps (<= MAX...: off)
		uint32_t rounds_last_frame = 0, readbacks_last_frame = 0;
	} m_rendering_buffer;

private:
	cudaEvent_t m_ev[3] = {nullptr, nullptr, nullptr};
public:
	bool m_shard_grid_refresh = true;  // world > 1: each rank refreshes 1/world of the occupancy samples, one allreduce(max) combines them (false: replicated)
	bool m_pipeline_march = true;      // false: issue the march on the main stream (serial schedule; used when timing kernels one by one)
private:
	std::chrono::steady_clock::time_point m_enqueue_start;
	cudaStream_t m_march_stream = nullptr;
	cudaStream_t m_band_stream[2] = {nullptr, nullptr};
	cudaEvent_t m_ev_frame_cleared = nullptr, m_ev_band_done[2] = {nullptr, nullptr};
	cudaEvent_t m_ev_march_go = nullptr, m_ev_march_ready[2] = {nullptr, nullptr}, m_ev_slot_free[2] = {nullptr, nullptr};
	void march_step(uint32_t slot, uint32_t target_batch_size, uint32_t n_rays_per_batch, uint32_t n_rays_total, default_rng_t rng, uint32_t* counters, cudaStream_t stream);
	void learn_step(uint32_t slot, uint32_t target_batch_size, uint32_t n_rays_per_batch, default_rng_t rng, uint32_t* counters, uint32_t* compacted_counter, float* loss,
	                cudaStream_t stream);
	uint32_t* m_host_counter = nullptr;    // pinned
	int m_dp_rank = 0, m_dp_world = 1;
	void* m_nccl_comm = nullptr;
	bool m_peer_enabled = false;
	ngp_peer_set m_peers{};
	DeviceBuffer<uint32_t> m_peer_flags;
	uint32_t m_peer_epoch = 0;
	std::vector<void*> m_peer_opened;
	bool m_dp_stale = false;                 // the modules were rebuilt after set_data_parallel: the peer set points at freed arenas
	bool m_sharded_state_dirty = false;      // peer mode: other ranks' shards of the optimizer state are older than the last step
	void close_peer_handles();
	int m_render_row_begin = 0, m_render_row_end = -1;
};

} // namespace ngp

// NCCL bootstrap for the data-parallel path (one process per GPU; the id travels through torch.distributed / any channel)
std::string ngp_host_nccl_unique_id();
void* ngp_host_nccl_comm_init(const std::string& id_bytes, int rank, int world_size);
void ngp_host_nccl_comm_destroy(void* comm);
