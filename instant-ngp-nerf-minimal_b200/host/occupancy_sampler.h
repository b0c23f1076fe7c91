// OccupancySampler -- same module boundary, method names and public fields as the reference's
// internal/sampler/occupancy_sampler.h:17-134; every method body is a call into the C ABI (include/ngp_b200.h).
#pragma once
#include "common.h"
#include "json.h"

namespace ngp {

struct NeRFSynthetic;

struct BoundingBox { float min[3], max[3]; };
struct SlotRange { uint32_t begin = 0, end = 0; };          // [begin, end) of the grid-sample arrays

class OccupancySampler {
public:
	explicit OccupancySampler(const Json& config);

	// occupancy_sampler.cu:633-658
	void init(uint32_t n_images, const int image_resolution[2], const float focal_length[2], const float principal_point[2],
	          const DeviceBuffer<Mat34>& camera_matrices, default_rng_t& rng, cudaStream_t stream);
	// :660-703
	// part / n_parts (data parallel): write only this rank's share; ranges[2] receives the slots written (uniform set, non-uniform set)
	void sample_positions_from_grid(uint32_t n_uniform_positions, uint32_t n_nonuniform_positions, DeviceBuffer<Position>& grid_positions,
	                                DeviceBuffer<uint32_t>& grid_indices, cudaStream_t stream, uint32_t part = 0, uint32_t n_parts = 1, SlotRange* ranges = nullptr);
	// :705-735
	void update_grid(uint32_t n_total_positions, uint32_t padded_output_width, const DeviceBuffer<precision_t>& grid_densities,
	                 const DeviceBuffer<uint32_t>& grid_indices, EActivation density_activation, cudaStream_t stream);
	// the two halves of update_grid, for an exchange (max over ranks of density_grid_current) between them
	void splat_grid(const SlotRange* ranges, uint32_t n_ranges, uint32_t padded_output_width, const DeviceBuffer<precision_t>& grid_densities,
	                const DeviceBuffer<uint32_t>& grid_indices, EActivation density_activation, cudaStream_t stream);
	void ema_grid(cudaStream_t stream);
	// :737-760
	void update_grid_bitfield(cudaStream_t stream);
	// :762-811.  scan_ws: scratch for the ordered compaction (sized by the callee)
	void sample_coords_from_dataset(uint32_t n_rays, uint32_t n_rays_shift, uint32_t n_coords, const NeRFSynthetic& dataset, bool train_envmap,
	                                DeviceBuffer<Ray>& rays, DeviceBuffer<float>& rays_rgba, uint32_t* n_rays_counter,
	                                DeviceBuffer<uint32_t>& ray_coord_indices, DeviceBuffer<Coordinate>& coords, uint32_t* n_coords_counter,
	                                default_rng_t rng, cudaStream_t stream);
	// :813-855
	void generate_rays_from_camera_matrix(const int resolution[2], const float focal_length[2], const float principal_point[2], const Mat34& camera_matrix,
	                                      uint32_t spp, DeviceBuffer<RayPayload>& payload, cudaStream_t stream,
	                                      int row_begin = 0, int row_end = -1,       // rows [row_begin, row_end) only; -1 = the whole image
	                                      const void* envmap = nullptr, const int* envmap_resolution = nullptr, float* frame_buffer = nullptr);   // :459-502
	// :857-877.  d_n_rays (optional): live-ray count in device memory, so no host round trip is needed
	void generate_coords_from_rays(uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps, DeviceBuffer<RayPayload>& payloads,
	                               DeviceBuffer<Coordinate>& coords, cudaStream_t stream);

	// device-driven render loop: live-ray count and step count come from the control block (include/ngp_b200.h)
	void generate_coords_from_rays_dev(uint32_t n_rays_bound, const uint32_t* d_ctl, uint32_t round, DeviceBuffer<RayPayload>& payloads,
	                                   DeviceBuffer<Coordinate>& coords, cudaStream_t stream);

	ngp_sampler_cfg cfg() const;

public:
	uint32_t aabb_scale;
	BoundingBox aabb;
	uint32_t n_grid_size, n_grid_elements, n_cascades, n_total_elements;
	float ema_decay;
	DeviceBuffer<float> density_grid_current, density_grid, density_grid_mean;
	DeviceBuffer<uint8_t> density_grid_bitfield;
	DeviceBuffer<uint8_t> density_grid_coarse;       // one bit per 8x8x8 block of cells (ngp_build_coarse_bitfield); empty when the grid is too small
	bool coarse_skip = true;                         // hand it to the inference march (identical results, fewer visited cells)
	uint32_t n_max_steps;
	float cone_angle_constant, min_cone_stepsize, max_cone_stepsize, near_distance, far_distance;
	bool snap_to_pixel_centers_in_training, snap_to_pixel_centers_in_rendering;
	float min_optical_thickness;
	uint32_t i_step = 0;
	default_rng_t rng;

private:
	DeviceBuffer<uint8_t> m_scan_ws;
};

} // namespace ngp
