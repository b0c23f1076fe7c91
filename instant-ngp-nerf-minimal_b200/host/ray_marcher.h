// RayMarcher -- the reference's internal/render/ray_marcher.h:18-85 module boundary over the C ABI compositing kernels.
#pragma once
#include "common.h"
#include "json.h"
#include "occupancy_sampler.h"

namespace ngp {

struct RaysSoA {   // internal/utils/common.h:146-153
	DeviceBuffer<RayPayload> payload;
	DeviceBuffer<float> rgba;          // 4 floats per ray
	void enlarge(size_t n) { payload.enlarge(n); rgba.enlarge(n * 4); }
};

class RayMarcher {
	const void* m_envmap = nullptr; void* m_envmap_gradient = nullptr; int m_envmap_res[2] = {0, 0}; ELossType m_envmap_loss_type = ELossType::L2;
public:
	explicit RayMarcher(const Json& config);

	// ray_marcher.cu:305-377.  n_rays_counter / n_compacted_coords_counter are device counters.
	void volume_rendering_with_loss_and_gradient(
		uint32_t n_rays, uint32_t n_max_compacted_coords,
		const DeviceBuffer<Ray>& rays, const DeviceBuffer<float>& rays_rgba, const uint32_t* n_rays_counter, DeviceBuffer<uint32_t>& ray_coord_indices,
		const DeviceBuffer<Coordinate>& coords, DeviceBuffer<Coordinate>& compacted_coords, uint32_t* n_compacted_coords_counter,
		const DeviceBuffer<precision_t>& network_output, DeviceBuffer<precision_t>& dloss_doutput, int padded_output_width,
		EActivation rgb_activation, EActivation density_activation,
		float* loss_output, ELossType loss_type, float loss_scale,
		const BoundingBox& aabb, uint32_t n_cascades, float min_cone_stepsize, const DeviceBuffer<float>& density_grid_mean, float min_optical_thickness,
		const float background_color[3], EColorSpace color_space, default_rng_t rng, cudaStream_t stream, uint32_t* compacted_src_index = nullptr);

	// data parallel: 1 / world, so the density L1 regulariser (added outside the loss scale) is summed once over the ranks
	void set_regulariser_scale(float s) { m_reg_scale = s; }
	float m_reg_scale = 1.f;
	// environment map arguments of ray_marcher.cu:43-46 (data, gradient, resolution, loss type); nullptr data = none
	void set_envmap(const void* data, void* gradient, int res_x, int res_y, ELossType loss_type) {
		m_envmap = data; m_envmap_gradient = gradient; m_envmap_res[0] = res_x; m_envmap_res[1] = res_y; m_envmap_loss_type = loss_type;
	}

	// ray_marcher.cu:379-416
	void volume_rendering(uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps, uint32_t i_steps_offset, RaysSoA& rays, const DeviceBuffer<Coordinate>& coords,
	                      const DeviceBuffer<precision_t>& rgbsigma, uint32_t padded_output_width, EActivation density_activation, EActivation rgb_activation,
	                      ERenderMode render_mode, const BoundingBox& aabb, uint32_t n_cascades, float min_cone_stepsize, const Mat34& camera_matrix,
	                      float depth_scale, cudaStream_t stream);
	// the same with the round's counts read from the device control block of the device-driven render loop (include/ngp_b200.h)
	// next / hit (optional, together): the launch also compacts -- live rays go to `next`, finished ones to the hit list -- so the
	// following round needs no compaction launch (ngp_composite_infer_compact_dev)
	void volume_rendering_dev(uint32_t n_rays_bound, uint32_t* d_ctl, uint32_t round, RaysSoA& rays, const DeviceBuffer<Coordinate>& coords,
	                          const DeviceBuffer<precision_t>& rgbsigma, uint32_t padded_output_width, EActivation density_activation, EActivation rgb_activation,
	                          ERenderMode render_mode, const BoundingBox& aabb, uint32_t n_cascades, float min_cone_stepsize, const Mat34& camera_matrix,
	                          float depth_scale, cudaStream_t stream, RaysSoA* next = nullptr, RaysSoA* hit = nullptr, uint32_t* d_n_hit = nullptr);

public:
	bool train_with_random_bg_color;
	bool train_in_linear_color;
	float transmittance_threshold;

private:
	DeviceBuffer<uint8_t> m_scan_ws;
};

} // namespace ngp
