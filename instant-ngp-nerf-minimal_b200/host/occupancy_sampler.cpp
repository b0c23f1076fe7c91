// OccupancySampler host methods: parameter bookkeeping identical to internal/sampler/occupancy_sampler.cu:603-877,
// kernels behind the C ABI.
#include "occupancy_sampler.h"
#include "dataset.h"

namespace ngp {

OccupancySampler::OccupancySampler(const Json& config) {
	aabb_scale = (uint32_t)config.value("aabb_scale", 1);
	for (int d = 0; d < 3; ++d) { aabb.min[d] = 0.5f - 0.5f * (float)aabb_scale; aabb.max[d] = 0.5f + 0.5f * (float)aabb_scale; }
	n_grid_size = (uint32_t)config.value("grid_size", 128);
	n_grid_elements = n_grid_size * n_grid_size * n_grid_size;
	n_cascades = 2;
	while ((1u << (n_cascades - 1)) < aabb_scale) n_cascades++;
	n_total_elements = n_grid_elements * n_cascades;
	ema_decay = config.value("ema_decay", 0.95f);
	density_grid_current.resize(n_total_elements);
	density_grid.resize(n_total_elements);
	density_grid_mean.resize(256);
	density_grid_bitfield.resize(n_total_elements / 8);
	density_grid_current.memset(0); density_grid.memset(0); density_grid_mean.memset(0); density_grid_bitfield.memset(0);
	if ((n_grid_size & (n_grid_size - 1)) == 0 && n_grid_size >= 32) {
		density_grid_coarse.resize(n_total_elements / 4096);
		density_grid_coarse.memset(0);
	}
	n_max_steps = (uint32_t)config.value("maximum_marching_steps", 1024);
	cone_angle_constant = config.value("cone_angle_constant", 0.f);
	const float sqrt3 = 1.73205080757f;
	min_cone_stepsize = sqrt3 / (float)n_max_steps;
	max_cone_stepsize = sqrt3 * (float)(1u << (n_cascades - 1)) / (float)n_grid_size;
	near_distance = config.value("near_distance", 0.2f);
	far_distance = config.value("far_distance", 100.f);
	snap_to_pixel_centers_in_training = config.value("snap_to_pixel_centers_in_training", true);
	snap_to_pixel_centers_in_rendering = config.value("snap_to_pixel_centers_in_rendering", true);
	min_optical_thickness = config.value("min_optical_thickness", 0.01f);
}

ngp_sampler_cfg OccupancySampler::cfg() const {
	ngp_sampler_cfg c{};
	c.aabb_min = aabb.min[0]; c.aabb_max = aabb.max[0];
	c.grid_size = n_grid_size; c.n_cascades = n_cascades; c.max_steps = n_max_steps;
	c.cone_angle = cone_angle_constant; c.min_step = min_cone_stepsize; c.max_step = max_cone_stepsize; c.near_distance = near_distance;
	c.coarse_bitfield = coarse_skip && density_grid_coarse.size() ? density_grid_coarse.data() : nullptr;
	return c;
}

void OccupancySampler::init(uint32_t n_images, const int image_resolution[2], const float focal_length[2], const float principal_point[2],
                            const DeviceBuffer<Mat34>& camera_matrices, default_rng_t& rng_in, cudaStream_t stream) {
	i_step = 0;
	rng = default_rng_t{(uint64_t)rng_in.next_uint()};
	const ngp_sampler_cfg c = cfg();
	ngp_camera_cfg cam{image_resolution[0], image_resolution[1], focal_length[0], focal_length[1], principal_point[0], principal_point[1]};
	api_check(ngp_mark_untrained_grid(stream, &c, &cam, n_images, (const float*)camera_matrices.data(), density_grid.data()));
}

void OccupancySampler::sample_positions_from_grid(uint32_t n_uniform_positions, uint32_t n_nonuniform_positions, DeviceBuffer<Position>& grid_positions,
                                                  DeviceBuffer<uint32_t>& grid_indices, cudaStream_t stream, uint32_t part, uint32_t n_parts, SlotRange* ranges) {
	const uint32_t n_total = n_uniform_positions + n_nonuniform_positions;
	grid_positions.enlarge(n_total);
	grid_indices.enlarge(n_total);
	const ngp_sampler_cfg c = cfg();
	SlotRange r[2] = {};
	// Morton-ordered emission: the same samples as the reference's kernel, in an order the hash-grid gathers like (see ngp_b200.h).
	// With n_parts > 1 this rank writes (and later evaluates) only its contiguous share of each of the two sample sets.
	api_check(ngp_generate_grid_samples_sorted_part(stream, &c, n_uniform_positions, rng.state, rng.inc, i_step, density_grid.data(), -0.01f,
	                                                (float*)grid_positions.data(), grid_indices.data(), part, n_parts, &r[0].begin, &r[0].end));
	rng.advance();
	api_check(ngp_generate_grid_samples_sorted_part(stream, &c, n_nonuniform_positions, rng.state, rng.inc, i_step, density_grid.data(), min_optical_thickness,
	                                                (float*)(grid_positions.data() + n_uniform_positions), grid_indices.data() + n_uniform_positions,
	                                                part, n_parts, &r[1].begin, &r[1].end));
	rng.advance();
	r[1].begin += n_uniform_positions; r[1].end += n_uniform_positions;
	if (ranges) { ranges[0] = r[0]; ranges[1] = r[1]; }
}

void OccupancySampler::update_grid(uint32_t n_total_positions, uint32_t padded_output_width, const DeviceBuffer<precision_t>& grid_densities,
                                   const DeviceBuffer<uint32_t>& grid_indices, EActivation density_activation, cudaStream_t stream) {
	const ngp_sampler_cfg c = cfg();
	api_check(ngp_update_grid(stream, &c, n_total_positions, grid_indices.data(), grid_densities.data(), padded_output_width, (int)density_activation,
	                          ema_decay, density_grid_current.data(), density_grid.data()));
	++i_step;
}

// update_grid in its two halves (data parallel: the ranks' partial maxima are combined between them)
void OccupancySampler::splat_grid(const SlotRange* ranges, uint32_t n_ranges, uint32_t padded_output_width, const DeviceBuffer<precision_t>& grid_densities,
                                  const DeviceBuffer<uint32_t>& grid_indices, EActivation density_activation, cudaStream_t stream) {
	const ngp_sampler_cfg c = cfg();
	density_grid_current.memset(0, stream);
	for (uint32_t k = 0; k < n_ranges; ++k)
		if (ranges[k].end > ranges[k].begin)
			api_check(ngp_splat_grid(stream, &c, ranges[k].end - ranges[k].begin, grid_indices.data() + ranges[k].begin,
			                         grid_densities.data() + (size_t)ranges[k].begin * padded_output_width, padded_output_width, (int)density_activation,
			                         density_grid_current.data()));
}

void OccupancySampler::ema_grid(cudaStream_t stream) {
	const ngp_sampler_cfg c = cfg();
	api_check(ngp_ema_grid(stream, &c, ema_decay, density_grid_current.data(), density_grid.data()));
	++i_step;
}

void OccupancySampler::update_grid_bitfield(cudaStream_t stream) {
	const ngp_sampler_cfg c = cfg();
	api_check(ngp_update_grid_bitfield(stream, &c, density_grid.data(), min_optical_thickness, density_grid_mean.data(), density_grid_bitfield.data()));
	if (density_grid_coarse.size()) api_check(ngp_build_coarse_bitfield(stream, &c, density_grid_bitfield.data(), density_grid_coarse.data()));
}

void OccupancySampler::sample_coords_from_dataset(uint32_t n_rays, uint32_t n_rays_shift, uint32_t n_coords, const NeRFSynthetic& dataset, bool train_envmap,
                                                  DeviceBuffer<Ray>& rays, DeviceBuffer<float>& rays_rgba, uint32_t* n_rays_counter,
                                                  DeviceBuffer<uint32_t>& ray_coord_indices, DeviceBuffer<Coordinate>& coords, uint32_t* n_coords_counter,
                                                  default_rng_t rng_step, cudaStream_t stream) {
	// the C ABI keeps the two counters adjacent: [0] rays kept, [1] samples counted
	if (n_coords_counter != n_rays_counter + 1) throw std::runtime_error("sample_coords_from_dataset: counters must be adjacent (rays, samples)");
	m_scan_ws.enlarge(ngp_scan_workspace_bytes(n_rays));
	ngp_sampler_cfg c = cfg();
	c.keep_empty_rays = train_envmap ? 1u : 0u;          // occupancy_sampler.cu:408
	ngp_camera_cfg cam{dataset.image_resolution[0], dataset.image_resolution[1], dataset.focal_length[0], dataset.focal_length[1],
	                   dataset.principal_point[0], dataset.principal_point[1]};
	api_check(ngp_generate_training_samples(stream, &c, &cam, n_rays, n_rays_shift, n_coords, dataset.n_images, (const float*)dataset.xforms.data(),
	                                        dataset.images.data(), density_grid_bitfield.data(), snap_to_pixel_centers_in_training ? 1 : 0,
	                                        rng_step.state, rng_step.inc, (float*)rays.data(), rays_rgba.data(), ray_coord_indices.data(),
	                                        (float*)coords.data(), n_rays_counter, nullptr, m_scan_ws.data()));
}

void OccupancySampler::generate_rays_from_camera_matrix(const int resolution[2], const float focal_length[2], const float principal_point[2],
                                                        const Mat34& camera_matrix, uint32_t spp, DeviceBuffer<RayPayload>& payload, cudaStream_t stream,
                                                        int row_begin, int row_end, const void* envmap, const int* envmap_resolution, float* frame_buffer) {
	const ngp_sampler_cfg c = cfg();
	ngp_camera_cfg cam{resolution[0], resolution[1], focal_length[0], focal_length[1], principal_point[0], principal_point[1]};
	if (row_end < 0) { row_begin = 0; row_end = resolution[1]; }
	if (envmap) {
		api_check(ngp_init_rays_rows_envmap(stream, &c, &cam, camera_matrix.m, spp, snap_to_pixel_centers_in_rendering ? 1 : 0, density_grid_bitfield.data(), payload.data(),
		                                    (uint32_t)row_begin, (uint32_t)row_end, envmap, envmap_resolution[0], envmap_resolution[1], frame_buffer));
		return;
	}
	api_check(ngp_init_rays_rows(stream, &c, &cam, camera_matrix.m, spp, snap_to_pixel_centers_in_rendering ? 1 : 0, density_grid_bitfield.data(), payload.data(),
	                             (uint32_t)row_begin, (uint32_t)row_end));
}

void OccupancySampler::generate_coords_from_rays(uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps, DeviceBuffer<RayPayload>& payloads,
                                                 DeviceBuffer<Coordinate>& coords, cudaStream_t stream) {
	const ngp_sampler_cfg c = cfg();
	api_check(ngp_generate_coords_from_rays(stream, &c, n_rays, d_n_rays, n_steps, density_grid_bitfield.data(), payloads.data(), (float*)coords.data()));
}

void OccupancySampler::generate_coords_from_rays_dev(uint32_t n_rays_bound, const uint32_t* d_ctl, uint32_t round, DeviceBuffer<RayPayload>& payloads,
                                                     DeviceBuffer<Coordinate>& coords, cudaStream_t stream) {
	const ngp_sampler_cfg c = cfg();
	api_check(ngp_generate_coords_from_rays_dev(stream, &c, n_rays_bound, d_ctl, round, density_grid_bitfield.data(), payloads.data(), (float*)coords.data()));
}

} // namespace ngp
