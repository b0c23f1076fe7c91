#include "ray_marcher.h"

namespace ngp {

RayMarcher::RayMarcher(const Json& config) {
	train_with_random_bg_color = config.value("train_with_random_bg_color", true);
	train_in_linear_color = config.value("train_in_linear_color", false);
	transmittance_threshold = config.value("transmittance_threshold", 0.0001f);
}

void RayMarcher::volume_rendering_with_loss_and_gradient(
	uint32_t n_rays, uint32_t n_max_compacted_coords,
	const DeviceBuffer<Ray>& rays, const DeviceBuffer<float>& rays_rgba, const uint32_t* n_rays_counter, DeviceBuffer<uint32_t>& ray_coord_indices,
	const DeviceBuffer<Coordinate>& coords, DeviceBuffer<Coordinate>& compacted_coords, uint32_t* n_compacted_coords_counter,
	const DeviceBuffer<precision_t>& network_output, DeviceBuffer<precision_t>& dloss_doutput, int padded_output_width,
	EActivation rgb_activation, EActivation density_activation,
	float* loss_output, ELossType loss_type, float loss_scale,
	const BoundingBox&, uint32_t n_cascades, float min_cone_stepsize, const DeviceBuffer<float>& density_grid_mean, float min_optical_thickness,
	const float background_color[3], EColorSpace color_space, default_rng_t rng, cudaStream_t stream, uint32_t* compacted_src_index)
{
	ngp_render_cfg c{};
	c.n_cascades = n_cascades; c.min_step = min_cone_stepsize;
	c.rgb_activation = (int)rgb_activation; c.density_activation = (int)density_activation;
	c.loss_type = (int)loss_type; c.color_space = (int)color_space;
	c.train_with_random_bg_color = train_with_random_bg_color ? 1 : 0; c.train_in_linear_color = train_in_linear_color ? 1 : 0;
	c.transmittance_threshold = transmittance_threshold; c.min_optical_thickness = min_optical_thickness; c.loss_scale = loss_scale;
	c.reg_scale = m_reg_scale;
	for (int k = 0; k < 3; ++k) c.background_color[k] = background_color[k];
	if (m_envmap) {
		c.envmap = m_envmap; c.envmap_gradient = m_envmap_gradient; c.rays = (const float*)rays.data();
		c.envmap_res_x = m_envmap_res[0]; c.envmap_res_y = m_envmap_res[1]; c.envmap_loss_type = (int)m_envmap_loss_type;
	}
	m_scan_ws.enlarge(ngp_composite_workspace_bytes(n_rays));
	api_check(ngp_composite_train(stream, &c, n_rays, n_max_compacted_coords, n_rays_counter, rays_rgba.data(), ray_coord_indices.data(),
	                              (const float*)coords.data(), (float*)compacted_coords.data(), n_compacted_coords_counter, network_output.data(),
	                              dloss_doutput.data(), (uint32_t)padded_output_width, density_grid_mean.data(), rng.state, rng.inc, loss_output, m_scan_ws.data(), compacted_src_index));
}

void RayMarcher::volume_rendering(uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps, uint32_t i_steps_offset, RaysSoA& rays, const DeviceBuffer<Coordinate>& coords,
                                  const DeviceBuffer<precision_t>& rgbsigma, uint32_t padded_output_width, EActivation density_activation, EActivation rgb_activation,
                                  ERenderMode render_mode, const BoundingBox& aabb, uint32_t n_cascades, float min_cone_stepsize, const Mat34& camera_matrix,
                                  float depth_scale, cudaStream_t stream)
{
	ngp_render_cfg c{};
	c.n_cascades = n_cascades; c.min_step = min_cone_stepsize;
	c.rgb_activation = (int)rgb_activation; c.density_activation = (int)density_activation;
	c.transmittance_threshold = transmittance_threshold;
	const float fwd[3] = {camera_matrix.at(0, 2), camera_matrix.at(1, 2), camera_matrix.at(2, 2)};
	api_check(ngp_composite_infer(stream, &c, n_rays, d_n_rays, n_steps, i_steps_offset, rays.payload.data(), rays.rgba.data(), (const float*)coords.data(),
	                              rgbsigma.data(), padded_output_width, (int)render_mode, aabb.min[0], aabb.max[0], fwd, depth_scale));
}

void RayMarcher::volume_rendering_dev(uint32_t n_rays_bound, uint32_t* d_ctl, uint32_t round, RaysSoA& rays, const DeviceBuffer<Coordinate>& coords,
                                      const DeviceBuffer<precision_t>& rgbsigma, uint32_t padded_output_width, EActivation density_activation, EActivation rgb_activation,
                                      ERenderMode render_mode, const BoundingBox& aabb, uint32_t n_cascades, float min_cone_stepsize, const Mat34& camera_matrix,
                                      float depth_scale, cudaStream_t stream, RaysSoA* next, RaysSoA* hit, uint32_t* d_n_hit)
{
	ngp_render_cfg c{};
	c.n_cascades = n_cascades; c.min_step = min_cone_stepsize;
	c.rgb_activation = (int)rgb_activation; c.density_activation = (int)density_activation;
	c.transmittance_threshold = transmittance_threshold;
	const float fwd[3] = {camera_matrix.at(0, 2), camera_matrix.at(1, 2), camera_matrix.at(2, 2)};
	if (next && hit && d_n_hit) {
		api_check(ngp_composite_infer_compact_dev(stream, &c, n_rays_bound, d_ctl, round, rays.payload.data(), rays.rgba.data(), (const float*)coords.data(),
		                                          rgbsigma.data(), padded_output_width, (int)render_mode, aabb.min[0], aabb.max[0], fwd, depth_scale,
		                                          next->payload.data(), next->rgba.data(), hit->payload.data(), hit->rgba.data(), d_n_hit));
		return;
	}
	api_check(ngp_composite_infer_dev(stream, &c, n_rays_bound, d_ctl, round, rays.payload.data(), rays.rgba.data(), (const float*)coords.data(),
	                                  rgbsigma.data(), padded_output_width, (int)render_mode, aabb.min[0], aabb.max[0], fwd, depth_scale));
}

} // namespace ngp
