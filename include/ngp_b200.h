/* ngp_b200.h -- C ABI of the B200-native instant-NGP NeRF hot path.
 *
 * Drop-in boundary for the reference's internal/{sampler,network,render} modules and the tiny-cuda-nn
 * objects they call (file:line below are relative to the reference tree).  Plain pointers and sizes only:
 * every pointer named d_* is DEVICE memory owned by the caller, every call is asynchronous on `stream`
 * (a cudaStream_t passed as void*), nothing allocates, nothing synchronises with the host (one set-up call,
 * ngp_grid_desc_device_scales, waits for its own 128-byte result and says so).
 * Return value: 0 on success, <0 on error (message via ngp_last_error()).  There is no CPU fallback.
 */
#ifndef NGP_B200_H
#define NGP_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NGP_MAX_LEVELS 32
#define NGP_MLP_PARAMS 10240u   /* W0[64][32] W1[16][64] V0[64][32] V1[64][64] V2[16][64], row-major [out][in] */

/* Multiresolution hash-grid geometry (tiny-cuda-nn/encodings/grid.h:475-537). Built by ngp_grid_desc_init. */
typedef struct ngp_grid_desc {
	uint32_t n_levels;
	uint32_t offsets[NGP_MAX_LEVELS + 1];  /* prefix sums of per-level entry counts (2 fp16 features / entry) */
	float    scales[NGP_MAX_LEVELS];       /* exp2f(l*log2(per_level_scale))*base - 1, host float math */
	uint32_t resolutions[NGP_MAX_LEVELS];  /* ceil(scale)+1 */
} ngp_grid_desc;

/* Occupancy sampler parameters (internal/sampler/occupancy_sampler.cu:603-631). */
typedef struct ngp_sampler_cfg {
	float    aabb_min, aabb_max;     /* cube [min,max]^3 = 0.5 -/+ 0.5*aabb_scale */
	uint32_t grid_size;              /* 128 */
	uint32_t n_cascades;             /* >= 2 */
	uint32_t max_steps;              /* maximum_marching_steps, 1024 */
	float    cone_angle;             /* cone_angle_constant */
	float    min_step, max_step;     /* sqrt3/max_steps, sqrt3*2^(C-1)/grid_size */
	float    near_distance;          /* 0.2 */
	uint32_t keep_empty_rays;        /* train_envmap (occupancy_sampler.cu:408): training rays without a single sample stay in the ray list */
	const uint8_t* coarse_bitfield;  /* optional (may be NULL): one bit per 8x8x8 block of cells, set when any cell of the block is occupied
	                                    (ngp_build_coarse_bitfield).  The inference march crosses an empty block in one hop when it can
	                                    prove the landing point is the one the cell-by-cell march reaches; results are identical. */
} ngp_sampler_cfg;

/* Pinhole camera set (internal/dataset/nerf_synthetic.h; xforms 3x4 column-major, 12 floats each). */
typedef struct ngp_camera_cfg {
	int32_t width, height;
	float   fx, fy, ppx, ppy;
} ngp_camera_cfg;

/* Compositing parameters (internal/render/ray_marcher.cu:15-54, internal/utils/common.h:57-94). */
typedef struct ngp_render_cfg {
	uint32_t n_cascades;
	float    min_step;
	int32_t  rgb_activation, density_activation;   /* EActivation: 0 None 1 ReLU 2 Logistic 3 Exponential */
	int32_t  loss_type;                            /* ELossType: 0 L2 1 L1 2 Mape 3 Smape 4 SmoothL1 5 LogL1 6 RelativeL2 */
	int32_t  color_space;                          /* EColorSpace: 0 Linear 1 SRGB */
	int32_t  train_with_random_bg_color, train_in_linear_color;
	float    transmittance_threshold;
	float    min_optical_thickness;
	float    loss_scale;                           /* 128 */
	float    background_color[3];
	/* trainable environment map (internal/utils/envmap.h, ray_marcher.cu:96-102,190-212); all zero / NULL = none.
	 * envmap: fp16 [res_y][res_x][4]; the reference fills the resolution as {height, width} and indexes x + y * res_x */
	const void* envmap;              /* read behind the volume: background = envmap.rgb + background * (1 - envmap.a) */
	void*       envmap_gradient;     /* fp16, same shape: dL/d(envmap.rgb) deposited bilinearly for fully traversed rays */
	const float* rays;               /* device, 6 floats per kept ray {o, d}: the directions the envmap is looked up with */
	int32_t  envmap_res_x, envmap_res_y;
	int32_t  envmap_loss_type;       /* ELossType of the envmap's own loss (base.json: RelativeL2) */
	/* scale of the density L1 regulariser (ray_marcher.cu:153,182), which the reference adds OUTSIDE the loss scale: data-parallel
	 * ranks pass 1 / world so that the summed gradient carries it once, like the loss terms.  0 is read as 1. */
	float    reg_scale;
} ngp_render_cfg;

/* Optimizer chain Ema(ExponentialDecay(Adam)) (tiny-cuda-nn/optimizers/{adam,ema,exponential_decay}.h). */
typedef struct ngp_optimizer_cfg {
	float learning_rate, beta1, beta2, epsilon, l2_reg, loss_scale, ema_decay;
} ngp_optimizer_cfg;

const char* ngp_last_error(void);
int ngp_abi_version(void);
/* measurement hooks: kernels launched by this library so far; optional CUDA-event bracketing of every launch on its own
 * stream (enable with a capacity, collect = synchronise + per-kernel-name totals, returns the number of names written) */
unsigned long long ngp_launch_count(void);
int ngp_kernel_timing_enable(int on, uint32_t max_launches);
int ngp_kernel_timing_collect(uint32_t max_kernels, char* names /* [max_kernels][64] */, float* total_ms, uint32_t* counts);
/* per-launch start / end (ms since the first recorded launch), in launch order; synchronises and resets like _collect */
int ngp_kernel_timing_timeline(uint32_t max_launches, char* names /* [max_launches][64] */, float* start_ms, float* end_ms);

/* Measured ceilings of the L1TEX / L2 path for the hash-grid kernels' access patterns ("L2 peak ... must be measured
 * on the box"): kind 0 = coalesced 16-byte reads of an L2-resident buffer (result in GB/s); 1 / 2 = scattered 4- / 16-byte gathers,
 * 3 / 4 = scattered 4- / 16-byte fp16 reductions, every lane its own line (result in 1e9 accesses / s).  d_table: table_bytes of
 * scratch (power of two; 32 MB stays L2-resident), zero-filled by the call.  Times `repeats` launches with CUDA events; synchronises. */
int ngp_l2_probe(void* stream, int kind, void* d_table, uint64_t table_bytes, uint32_t repeats, double* result);

/* ---- hash grid -------------------------------------------------------------------------------- */
/* replaces GridEncodingTemplated ctor (grid.h:475-537); returns the number of table entries */
uint32_t ngp_grid_desc_init(ngp_grid_desc* desc, uint32_t n_levels, uint32_t log2_hashmap_size, uint32_t base_resolution, float per_level_scale);
/* replaces desc->scales / resolutions by what the reference's kernels evaluate per thread ON THE DEVICE (grid.h:157-158: exp2f ->
 * MUFU.EX2, one FFMA): an ulp away from the host value on some levels.  Offsets are untouched (the reference sizes the levels on
 * the host, grid.h:497-510).  Set-up call: launches one tiny kernel on `stream` into d_scratch (NGP_MAX_LEVELS floats of device
 * memory owned by the caller), copies the result back and WAITS for it (the one entry point that synchronises `stream`). */
int ngp_grid_desc_device_scales(ngp_grid_desc* desc, uint32_t base_resolution, float per_level_scale, void* stream, float* d_scratch);
/* replaces NGPNetwork::auto_determine_grid_parameters (internal/network/ngp_network.h:597-641) */
float ngp_auto_per_level_scale(float desired_resolution, uint32_t base_resolution, uint32_t n_levels);

/* replaces GridEncodingTemplated::encode (grid.h:539-610): d_out[i*out_stride + 2*l + f], fp16.
 * d_indices (optional, tests): [n][n_levels][8] u32 entry index of each corner. */
int ngp_hash_encode_fwd(void* stream, const float* d_xyz, uint32_t stride_floats, uint32_t n, const void* d_table,
                        const ngp_grid_desc* desc, void* d_out, uint32_t out_stride, uint32_t* d_indices);
/* replaces GridEncodingTemplated::backward (grid.h:612-696): zeroes d_grad (fp16, 2*entries) when zero_first,
 * then scatters w*dL/dy with warp-aggregated half2 atomics. */
int ngp_hash_encode_bwd(void* stream, const float* d_xyz, uint32_t stride_floats, uint32_t n, const void* d_dy, uint32_t dy_stride,
                        const ngp_grid_desc* desc, void* d_grad, int zero_first);

/* ---- network ---------------------------------------------------------------------------------- */
/* d_params: fp16 arena [NGP_MLP_PARAMS | grid table], reference order (ngp_network.h:496-541). */
size_t ngp_network_workspace_bytes(uint32_t n);
/* replaces NGPNetwork::inference_mixed_precision (ngp_network.h:217-284): Coordinate[n] (7 floats) -> rgbsigma [n][4] fp16.
 * d_n_valid (optional): device counter; tiles beyond it are skipped.  On return the first n*64 bytes of d_workspace hold
 * the encoded positions [n][32] fp16. */
int ngp_network_inference(void* stream, const float* d_coords, uint32_t n, const uint32_t* d_n_valid, const void* d_params,
                          const ngp_grid_desc* desc, void* d_rgbsigma, void* d_workspace, size_t workspace_bytes);
/* replaces NGPNetwork::density (ngp_network.h:286-307): positions (stride floats) -> [n][16] fp16 */
int ngp_network_density(void* stream, const float* d_positions, uint32_t stride_floats, uint32_t n, const void* d_params,
                        const ngp_grid_desc* desc, void* d_out, void* d_workspace, size_t workspace_bytes);
/* NGPNetwork::input_gradient(stream, 3, coords, coords) (tiny-cuda-nn/object.h:126-150; testbed.cu:558-560, RenderMode::Normals):
 * d sigma_raw / d input written OVER the 7-float Coordinates: position <- gradient, warped dt <- dt / 128 (the reference rescales
 * the row its backward pass never wrote), direction <- 0.  d_workspace as for ngp_network_train. */
int ngp_network_input_gradient(void* stream, float* d_coords, uint32_t n, const uint32_t* d_n_valid, const void* d_params,
                               const ngp_grid_desc* desc, void* d_workspace, size_t workspace_bytes);
/* replaces NGPNetwork::forward + ::backward (ngp_network.h:325-494) on the compacted batch: writes rgbsigma [n][4]
 * (optional) and the fp16 gradient arena d_grads [NGP_MLP_PARAMS | grid] (overwritten, not accumulated). */
int ngp_network_train(void* stream, const float* d_coords, uint32_t n, const void* d_params, const ngp_grid_desc* desc,
                      const void* d_dloss, void* d_rgbsigma, void* d_grads, void* d_workspace, size_t workspace_bytes);
/* Same result as ngp_network_train, without encoding the batch again: sample i takes row d_src_index[i] of d_encoded
 * ([m][32] fp16 -- the first m*64 bytes ngp_network_inference left in ITS workspace for the marched samples, of which the
 * compacted batch is a subset).  The table must not have changed in between.  d_workspace must not alias d_encoded. */
int ngp_network_train_reuse(void* stream, const float* d_coords, uint32_t n, const void* d_params, const ngp_grid_desc* desc,
                            const void* d_dloss, void* d_rgbsigma, void* d_grads, const void* d_encoded, uint32_t n_encoded /* = m; indices are clamped to it */,
                            const uint32_t* d_src_index, void* d_workspace, size_t workspace_bytes);
/* the two MLPs alone on a precomputed encoding (tests, profiling): enc [n][32] fp16 */
int ngp_mlp_inference(void* stream, const void* d_enc, const float* d_coords, uint32_t n, const void* d_mlp_params, void* d_rgbsigma);
int ngp_mlp_train(void* stream, const void* d_enc, const float* d_coords, uint32_t n, const void* d_mlp_params, const void* d_dloss,
                  void* d_rgbsigma, void* d_denc, float* d_wgrad_f32 /* [NGP_MLP_PARAMS], accumulated into */);

/* ---- sampler ---------------------------------------------------------------------------------- */
/* OccupancySampler::init -> mark_untrained_density_grid (occupancy_sampler.cu:18-61) */
int ngp_mark_untrained_grid(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, uint32_t n_images, const float* d_xforms, float* d_grid);
/* sample_positions_from_grid -> generate_grid_samples_nerf_nonuniform (:63-105) */
int ngp_generate_grid_samples(void* stream, const ngp_sampler_cfg* cfg, uint32_t n, uint64_t rng_state, uint64_t rng_inc, uint32_t step,
                              const float* d_grid, float thresh, float* d_positions, uint32_t* d_indices);
/* The same n samples -- the identical multiset of (position, index) pairs -- written in the Morton order of each sample's first
 * candidate cell instead of in sample order, so the density pass that follows reads spatially coherent positions.  Everything
 * downstream (density per position, max-splat by index) is order-free.  Needs grid_size^3 to be a power of two >= 256, else the
 * plain order is written. */
int ngp_generate_grid_samples_sorted(void* stream, const ngp_sampler_cfg* cfg, uint32_t n, uint64_t rng_state, uint64_t rng_inc, uint32_t step,
                                     const float* d_grid, float thresh, float* d_positions, uint32_t* d_indices);
/* One rank's share of ngp_generate_grid_samples_sorted under data parallelism: only the samples whose first candidate cell lies in
 * part `part` of `n_parts` contiguous Morton-cell ranges are written, into the SAME slots of d_positions / d_indices the full
 * call would use; *slot_begin / *slot_end (host) receive that contiguous slot range, so the density pass and the splat run on
 * [slot_begin, slot_end) alone.  The union over the parts is the full call's output, bit for bit. */
int ngp_generate_grid_samples_sorted_part(void* stream, const ngp_sampler_cfg* cfg, uint32_t n, uint64_t rng_state, uint64_t rng_inc, uint32_t step,
                                          const float* d_grid, float thresh, float* d_positions, uint32_t* d_indices,
                                          uint32_t part, uint32_t n_parts, uint32_t* slot_begin, uint32_t* slot_end);
/* update_grid (:705-735) in its two halves, for callers that put an exchange between them (max over ranks of d_grid_current):
 * splat_grid_samples_nerf_max_nearest_neighbor (:107-132) WITHOUT the memset, then ema_grid_samples_nerf (:134-158) */
int ngp_splat_grid(void* stream, const ngp_sampler_cfg* cfg, uint32_t n, const uint32_t* d_indices, const void* d_density, uint32_t density_stride,
                   int density_activation, float* d_grid_current);
int ngp_ema_grid(void* stream, const ngp_sampler_cfg* cfg, float decay, const float* d_grid_current, float* d_grid);
/* update_grid (:705-735): memset current, splat max, decay max */
int ngp_update_grid(void* stream, const ngp_sampler_cfg* cfg, uint32_t n, const uint32_t* d_indices, const void* d_density, uint32_t density_stride,
                    int density_activation, float decay, float* d_grid_current, float* d_grid);
/* update_grid_bitfield (:737-760): mean of cascade 0 -> d_mean[0] (d_mean: 1 result + 128 floats of scratch), bits, max-pool */
int ngp_update_grid_bitfield(void* stream, const ngp_sampler_cfg* cfg, const float* d_grid, float min_optical_thickness, float* d_mean, uint8_t* d_bitfield);
/* One bit per 8x8x8 block of cells (Morton order: block = cell index >> 9), cascades one after the other; d_coarse holds
 * ngp_coarse_bitfield_bytes(cfg) bytes.  grid_size must be a power of two >= 32.  No counterpart in the reference (which visits
 * every empty cell); see ngp_sampler_cfg.coarse_bitfield. */
size_t ngp_coarse_bitfield_bytes(const ngp_sampler_cfg* cfg);
int ngp_build_coarse_bitfield(void* stream, const ngp_sampler_cfg* cfg, const uint8_t* d_bitfield, uint8_t* d_coarse);
/* sample_coords_from_dataset -> generate_training_samples_nerf (:310-449).  Compaction is in ray order (deterministic).
 * d_counters: [0] rays kept, [1] samples counted (overwritten).  d_scan_ws: >= ngp_scan_workspace_bytes(n_rays) bytes of scratch. */
size_t ngp_scan_workspace_bytes(uint32_t n);
int ngp_generate_training_samples(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, uint32_t n_rays, uint32_t n_rays_shift,
                                  uint32_t n_coords, uint32_t n_images, const float* d_xforms, const void* d_images, const uint8_t* d_bitfield,
                                  int snap_to_pixel_centers, uint64_t rng_state, uint64_t rng_inc, float* d_rays, float* d_rays_rgba,
                                  uint32_t* d_numsteps, float* d_coords, uint32_t* d_counters, uint32_t* d_per_ray_n, void* d_scan_ws);
/* generate_rays_from_camera_matrix (:813-855): init_rays_with_payload_kernel_nerf + advance_pos_nerf; payload = 40 B records */
int ngp_init_rays(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, const float* camera_matrix_host, uint32_t spp,
                  int snap_to_pixel_centers, const uint8_t* d_bitfield, void* d_payloads);
/* the same for the rows [row_begin, row_end) of the image only (inference sharded over image tiles, one band per GPU, no
 * collective): payload slot i holds pixel row_begin * width + i, and everything derived from the pixel uses that global index,
 * so a band is bit-identical to the same rows of the full frame */
int ngp_init_rays_rows(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, const float* camera_matrix_host, uint32_t spp,
                       int snap_to_pixel_centers, const uint8_t* d_bitfield, void* d_payloads, uint32_t row_begin, uint32_t row_end);
/* with a trained environment map (occupancy_sampler.cu:459-502): additionally frame_buffer[pixel] <- read_envmap(ray direction) */
int ngp_init_rays_rows_envmap(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, const float* camera_matrix_host, uint32_t spp,
                              int snap_to_pixel_centers, const uint8_t* d_bitfield, void* d_payloads, uint32_t row_begin, uint32_t row_end,
                              const void* d_envmap, int32_t env_res_x, int32_t env_res_y, float* d_frame_buffer);
/* generate_coords_from_rays (:857-877) */
int ngp_generate_coords_from_rays(void* stream, const ngp_sampler_cfg* cfg, uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps,
                                  const uint8_t* d_bitfield, void* d_payloads, float* d_coords);

/* ---- render ----------------------------------------------------------------------------------- */
/* RayMarcher::volume_rendering_with_loss_and_gradient (ray_marcher.cu:15-213, 305-377), compaction in ray order.
 * d_scan_ws: >= ngp_composite_workspace_bytes(n_rays) bytes of scratch.
 * d_src_index (optional, [max_compacted]): for every compacted sample, its index in d_coords -- lets
 * ngp_network_train_reuse pick up the encodings ngp_network_inference already computed for it. */
size_t ngp_composite_workspace_bytes(uint32_t n_rays);
int ngp_composite_train(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays, uint32_t max_compacted, const uint32_t* d_ray_counter,
                        const float* d_rays_rgba, uint32_t* d_numsteps, const float* d_coords, float* d_coords_compacted,
                        uint32_t* d_compacted_counter, const void* d_rgbsigma, void* d_dloss, uint32_t padded_width,
                        const float* d_mean_density, uint64_t rng_state, uint64_t rng_inc, float* d_loss, void* d_scan_ws, uint32_t* d_src_index);
/* ngp_composite_train WITHOUT any allocation on either side: no counterpart in the reference -- added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Identical launches and results; the per-ray scratch (what d_scan_ws is above) lives in a file-scope `static __device__` array of
 * the library, sized for NGP_COMPOSITE_STATIC_MAX_RAYS rays (1.5 MB, held by every GPU that loads the library until exit).  Returns
 * -1 with a message when n_rays exceeds that maximum.  Calls no cudaMalloc / cudaFree / allocating routine.  The array is shared by
 * all calls and keeps the last call's contents (nothing in it has to start at zero); calls that overlap on one GPU -- two streams,
 * two threads -- race on it: serialise them, or use ngp_composite_train with separate workspaces. */
#define NGP_COMPOSITE_STATIC_MAX_RAYS 65536u
int ngp_composite_train_static(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays, uint32_t max_compacted, const uint32_t* d_ray_counter,
                               const float* d_rays_rgba, uint32_t* d_numsteps, const float* d_coords, float* d_coords_compacted,
                               uint32_t* d_compacted_counter, const void* d_rgbsigma, void* d_dloss, uint32_t padded_width,
                               const float* d_mean_density, uint64_t rng_state, uint64_t rng_inc, float* d_loss, uint32_t* d_src_index);
/* fill_rollover_and_rescale + fill_rollover (tiny-cuda-nn/common_device.h:485-505; testbed.cu:398-404) */
int ngp_rollover(void* stream, uint32_t batch, const uint32_t* d_compacted_counter, void* d_dloss, uint32_t padded_width, float* d_coords_compacted,
                 uint32_t* d_src_index /* optional, rolled over like the coordinates */);
/* RayMarcher::volume_rendering (ray_marcher.cu:215-295, 379-416); camera_fwd = camera matrix column 2 */
int ngp_composite_infer(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps, uint32_t n_steps_offset,
                        void* d_payloads, float* d_rgba, const float* d_coords, const void* d_rgbsigma, uint32_t padded_width, int render_mode,
                        float aabb_min, float aabb_max, const float* camera_fwd_host, float depth_scale);
/* Device-driven render loop.  Testbed::render_frame (testbed.cu:521-575) decides every round on the HOST how many steps to take
 * (n_steps = clamp(n_pixels / n_alive, MIN, MAX)) and so synchronises once per round; here the decision is a one-thread kernel and
 * the march / network / compositing launches read the round's counts from device memory, so a caller enqueues any number of rounds
 * and reads the live-ray count back when it likes.  d_ctl: NGP_RENDER_CTL_WORDS uint32, zeroed before round 0:
 *   [0],[1] live rays (ping-pong: the compaction of round r takes its input count from [r & 1] -- pass NULL in round 0 -- and must
 *           be given d_ctl + ((r + 1) & 1) as its live-ray counter),  [2] steps of the round,  [3] samples of the round
 *           (d_n_valid of ngp_network_inference),  [4] i_steps_offset.
 * Call order per round r: ngp_compact_rays, ngp_render_round, ngp_generate_coords_from_rays_dev, ngp_network_inference,
 * ngp_composite_infer_dev.  The step count is clamp(max(n_pixels, min_batch) / n_alive, min_steps, max_steps) -- min_batch = 0 is
 * the reference's rule; a positive value keeps small frames and row bands from taking one step per round --, and tail_steps >
 * max_steps lets late rounds (batch below a quarter of that) take up to tail_steps steps: fewer, fuller rounds.  A ray's samples
 * and their compositing order -- the image -- do not depend on the grouping.  Samples per round <= max(min(max(n_pixels, min_batch),
 * bound * max_steps), bound * min_steps, min(max(n_pixels, min_batch) / 4, bound * tail_steps)) for a live-ray bound `bound`. */
#define NGP_RENDER_CTL_WORDS 8
int ngp_render_round(void* stream, uint32_t* d_ctl, uint32_t round, uint32_t n_pixels, uint32_t min_steps, uint32_t max_steps, uint32_t tail_steps, uint32_t march_iter,
                     uint32_t min_batch);
int ngp_generate_coords_from_rays_dev(void* stream, const ngp_sampler_cfg* cfg, uint32_t n_rays_bound, const uint32_t* d_ctl, uint32_t round,
                                      const uint8_t* d_bitfield, void* d_payloads, float* d_coords);
int ngp_composite_infer_dev(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays_bound, const uint32_t* d_ctl, uint32_t round,
                            void* d_payloads, float* d_rgba, const float* d_coords, const void* d_rgbsigma, uint32_t padded_width, int render_mode,
                            float aabb_min, float aabb_max, const float* camera_fwd_host, float depth_scale);
/* ngp_composite_infer_dev followed by the compaction of the NEXT round in the same launch: after its compositing a live ray moves to
 * d_next_*, a finished one with alpha > 0.001 to the hit list, and the live count of round + 1 accumulates in d_ctl[round & 1] (the
 * counter ngp_render_round(round) has cleared) -- so only round 0 of a frame needs ngp_compact_rays. */
int ngp_composite_infer_compact_dev(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays_bound, uint32_t* d_ctl, uint32_t round,
                                    void* d_payloads, float* d_rgba, const float* d_coords, const void* d_rgbsigma, uint32_t padded_width, int render_mode,
                                    float aabb_min, float aabb_max, const float* camera_fwd_host, float depth_scale,
                                    void* d_next_payloads, float* d_next_rgba, void* d_hit_payloads, float* d_hit_rgba, uint32_t* d_n_hit);
/* compact_rays_kernel (testbed.cu:423-446) with warp-aggregated reservation; counters: [0] alive, [1] hit */
int ngp_compact_rays(void* stream, uint32_t n_rays, const uint32_t* d_n_rays, const void* d_src_payloads, const float* d_src_rgba,
                     void* d_dst_payloads, float* d_dst_rgba, void* d_hit_payloads, float* d_hit_rgba, uint32_t* d_n_alive, uint32_t* d_n_hit);
/* shade_kernel (testbed.cu:448-477) */
int ngp_shade(void* stream, uint32_t n_hit_max, const uint32_t* d_n_hit, const float* d_rgba, const void* d_payloads, float* d_frame, int render_mode, int train_in_linear);
/* CudaRenderBuffer::accumulate / ::tonemap (render_buffer.cu:51-186) on a linear float4 buffer */
int ngp_accumulate(void* stream, uint32_t n_pixels, const float* d_frame, float* d_accum, float sample_count, int color_space);
int ngp_tonemap(void* stream, uint32_t n_pixels, float exposure, const float* background_rgba_host, const float* d_accum, int color_space,
                int output_color_space, int tonemap_curve, float* d_out);

/* ---- parameters / optimizer ------------------------------------------------------------------- */
/* Trainer::initialize_params (trainer.h:73-112): xavier MLP (host draws) + grid U(-1e-4,1e-4); fills fp32 master and fp16 params.
 * rng_state is advanced exactly like the reference's host generator. */
int ngp_init_params(void* stream, uint64_t* rng_state, uint64_t rng_inc, uint64_t n_grid_params, float* d_params_fp32, void* d_params_fp16);
/* Trainer::optimizer_step (trainer.h:198) -> ExponentialDecay -> Adam -> EMA, one fused pass.  step = Adam step count BEFORE this step. */
int ngp_optimizer_step(void* stream, const ngp_optimizer_cfg* cfg, uint64_t n_params, uint64_t n_matrix_params, uint32_t step, float learning_rate,
                       float* d_w32, void* d_w16, const void* d_g16, float* d_m1, float* d_m2, uint32_t* d_param_steps, void* d_ema16);
/* ---- data-parallel optimizer over NVLink peer memory ------------------------------------------------
 * One process per GPU.  Instead of "all-reduce the gradient arena, then every rank runs the whole optimizer", rank r owns the
 * parameter range [shard_begin, shard_end): ONE kernel waits until every rank has published its gradients, sums the shard's
 * gradients straight out of the peers' arenas (P2P loads, fixed rank order, fp32), applies the same Adam + EMA update as
 * ngp_optimizer_step to the shard's state, and stores the new fp16 weights into every rank's parameter arena (P2P stores):
 * reduce-scatter + optimizer + all-gather without staging buffers, and 1/world of the optimizer work per GPU.
 * Pointers in ngp_peer_set are addresses valid ON THIS DEVICE (own allocations, and peers' opened through CUDA IPC).
 * flags: per rank, 2*NGP_MAX_PEERS uint32 in that rank's memory, zero-initialised; [p] = last epoch whose gradients rank p
 * published to me, [NGP_MAX_PEERS + p] = last epoch whose weights rank p finished storing into my arena. */
#define NGP_MAX_PEERS 8
typedef struct ngp_peer_set {
	uint32_t world, rank;
	const void* grads[NGP_MAX_PEERS];     /* fp16 gradient arena of rank p */
	void* params[NGP_MAX_PEERS];          /* fp16 parameter arena of rank p */
	uint32_t* flags[NGP_MAX_PEERS];       /* flag block of rank p */
} ngp_peer_set;
/* after the backward pass: tell every rank that this rank's gradients of `epoch` (>= 1, increasing) are complete */
int ngp_peer_publish_gradients(void* stream, const ngp_peer_set* peers, uint32_t epoch);
/* the fused step on [shard_begin, shard_end) (multiples of 8); state arrays are full-size, only the shard is touched */
int ngp_optimizer_step_sharded(void* stream, const ngp_optimizer_cfg* cfg, uint64_t shard_begin, uint64_t shard_end, uint64_t n_matrix_params, uint32_t step,
                               float learning_rate, float* d_w32, float* d_m1, float* d_m2, uint32_t* d_param_steps, void* d_ema16,
                               const ngp_peer_set* peers, uint32_t epoch, uint32_t* d_ticket /* one zeroed uint32 of scratch */);
/* before anything reads the parameters again: wait until every rank has stored its shard of `epoch` into this rank's arena */
int ngp_peer_wait_weights(void* stream, const ngp_peer_set* peers, uint32_t epoch);
/* free-running trace of the data-parallel step: host_out[0][e % 4096] = globaltimer ns at which this rank published the gradients
   of epoch e, host_out[1][e % 4096] = when all weight shards of epoch e had landed.  Synchronises the device. */
int ngp_peer_stamps_read(uint64_t* host_out /* [2][4096] */);

/* from_rgba32<__half> (internal/utils/common_device.h:160-175) */
int ngp_from_rgba32(void* stream, uint64_t n_pixels, const uint8_t* d_rgba8, void* d_out_half);
/* reduce_sum of the per-ray loss (testbed.cu:288).  d_out: 1 result float followed by >= 599 floats of scratch; the word
 * d_out[1] is a ticket that must be zero before the first use (the kernel resets it). */
int ngp_reduce_sum(void* stream, const float* d_in, uint64_t n, float* d_out);

/* ---- benchmark fixture ------------------------------------------------------------------------ */
/* Procedural stand-in for the Blender nerf_synthetic renders (no dataset can be downloaded on the GPU box): ray-marches an
 * analytic density / colour field inside the unit cube into n_views 8-bit sRGB RGBA images (straight alpha), which then go
 * through ngp_from_rgba32 like the reference loader's PNGs (internal/dataset/nerf_synthetic.cu:113-127).
 * d_xforms: n_views 3x4 column-major camera-to-world matrices in the unit-cube frame. */
#define NGP_SCENE_BLOBS 8
int ngp_render_procedural_views(void* stream, uint32_t seed, uint32_t n_views, int width, int height, float fx, float fy,
                                const float* d_xforms, uint8_t* d_rgba8);

#ifdef __cplusplus
}
#endif
#endif /* NGP_B200_H */
