// Volume-render compositing for training: forward, loss, analytic backward, ordered compaction.  sm_100a.
// Replaces volume_rendering_with_loss_and_gradient_kernal (internal/render/ray_marcher.cu:15-213) and the two rollover
// kernels (tiny-cuda-nn/common_device.h:485-505).  One thread walks one ray (rays are contiguous sample runs); the
// compacted base of each ray comes from the chained CTA scan in ray order instead of one global atomicAdd per ray, which
// makes the compacted batch bit-reproducible.  HBM-bound: 8 B rgbsigma + 28 B coord read, 8 B grad + 28 B coord written
// per sample.
#include "composite.cuh"

namespace ngp {

__global__ void __launch_bounds__(128) composite_train_kernel(
	ngp_render_cfg c, uint32_t n_rays, uint32_t max_compacted, const uint32_t* __restrict__ ray_counter,
	const float4* __restrict__ rays_rgba, uint32_t* __restrict__ numsteps_io, const Coordinate* __restrict__ coords_in,
	Coordinate* __restrict__ coords_compacted, uint32_t* __restrict__ compacted_counter, const __half* __restrict__ net_out,
	__half* __restrict__ dloss, uint32_t padded_width, const float* __restrict__ mean_density_ptr, Pcg32 rng,
	float* __restrict__ loss_out, ScanState scan)
{
	const uint32_t ticket = scan_ticket(scan);
	const uint32_t i = ticket * blockDim.x + threadIdx.x;
	const uint32_t n_kept = min(*ray_counter, n_rays);
	const bool active = i < n_kept;

	uint32_t numst = 0, base = 0, n_used = 0;
	float T = 1.f, rgb_ray[3] = {0.f, 0.f, 0.f};
	if (active) {
		numst = numsteps_io[i * 2]; base = numsteps_io[i * 2 + 1];
		const Coordinate* cs = coords_in + base;
		const __half* no = net_out + (size_t)base * 4;
		for (; n_used < numst; ++n_used) {
			if (T < c.transmittance_threshold) break;
			float raw[4];
			load_rgbsigma(no + (size_t)n_used * 4, raw);
			const float density = net_to_density(raw[3], c.density_activation);
			const float dt = unwarp_dt(cs[n_used].dt, c.min_step, c.n_cascades);
			const float alpha = 1.f - __expf(-density * dt), w = alpha * T;
			rgb_ray[0] += w * net_to_rgb(raw[0], c.rgb_activation);
			rgb_ray[1] += w * net_to_rgb(raw[1], c.rgb_activation);
			rgb_ray[2] += w * net_to_rgb(raw[2], c.rgb_activation);
			T *= (1.f - alpha);
		}
	}

	const unsigned long long prefix = chained_exclusive_scan(scan, ticket, (unsigned long long)n_used);
	const uint32_t cbase = (uint32_t)prefix;
	// final counter value (may exceed max_compacted, like the reference's atomicAdd total)
	if (i == (n_kept ? n_kept - 1 : 0) && (active || n_kept == 0)) *compacted_counter = n_kept ? cbase + n_used : 0;
	if (!active) return;

	// background: same random colour for every ray of a step (un-advanced generator, ray_marcher.cu:90-93)
	float bg[3] = {c.background_color[0], c.background_color[1], c.background_color[2]};
	if (c.train_with_random_bg_color) { bg[0] = rng.next_float(); bg[1] = rng.next_float(); bg[2] = rng.next_float(); }
	#pragma unroll
	for (int k = 0; k < 3; ++k) bg[k] = srgb_to_linear(bg[k]);
	const float4 gt4 = rays_rgba[i];
	const float gt[4] = {gt4.x, gt4.y, gt4.z, gt4.w};
	float tgt[3];
	if (c.train_in_linear_color || c.color_space == 0) {
		#pragma unroll
		for (int k = 0; k < 3; ++k) tgt[k] = gt[k] + (1 - gt[3]) * bg[k];
		if (!c.train_in_linear_color) {
			#pragma unroll
			for (int k = 0; k < 3; ++k) { tgt[k] = linear_to_srgb(tgt[k]); bg[k] = linear_to_srgb(bg[k]); }
		}
	} else {
		#pragma unroll
		for (int k = 0; k < 3; ++k) bg[k] = linear_to_srgb(bg[k]);
		#pragma unroll
		for (int k = 0; k < 3; ++k) tgt[k] = gt[3] > 0 ? linear_to_srgb(gt[k] / gt[3]) * gt[3] + (1 - gt[3]) * bg[k] : bg[k];
	}
	if (n_used == numst) {
		#pragma unroll
		for (int k = 0; k < 3; ++k) rgb_ray[k] += T * bg[k];
	}

	const uint32_t n = min(max_compacted - min(max_compacted, cbase), n_used);
	numsteps_io[i * 2] = n; numsteps_io[i * 2 + 1] = cbase;
	if (n == 0) return;

	float loss[3], g[3];
	loss_and_grad(c.loss_type, tgt, rgb_ray, loss, g);
	if (loss_out) loss_out[i] = ((loss[0] + loss[1] + loss[2]) / 3.f) / (float)n_rays;
	const float ls = c.loss_scale / (float)n_rays;
	const float l2reg = c.rgb_activation == ACT_EXP ? 1e-4f : 0.f;
	const float l1reg = *mean_density_ptr < c.min_optical_thickness ? 1e-4f : 0.f;

	const Coordinate* cs = coords_in + base;
	const __half* no = net_out + (size_t)base * 4;
	float rgb2[3] = {0.f, 0.f, 0.f};
	T = 1.f;
	for (uint32_t j = 0; j < n; ++j) {
		const Coordinate co = cs[j];
		coords_compacted[cbase + j] = co;
		float raw[4];
		load_rgbsigma(no + (size_t)j * 4, raw);
		const float dt = unwarp_dt(co.dt, c.min_step, c.n_cascades);
		const float rgb[3] = {net_to_rgb(raw[0], c.rgb_activation), net_to_rgb(raw[1], c.rgb_activation), net_to_rgb(raw[2], c.rgb_activation)};
		const float density = net_to_density(raw[3], c.density_activation);
		const float alpha = 1.f - __expf(-density * dt), w = alpha * T;
		#pragma unroll
		for (int k = 0; k < 3; ++k) rgb2[k] += w * rgb[k];
		T *= (1.f - alpha);
		float dl[4];
		#pragma unroll
		for (int k = 0; k < 3; ++k) dl[k] = ls * (w * g[k] * net_to_rgb_d(raw[k], c.rgb_activation) + fmaxf(0.f, l2reg * raw[k]));
		float dot = 0.f;
		#pragma unroll
		for (int k = 0; k < 3; ++k) dot += g[k] * (T * rgb[k] - (rgb_ray[k] - rgb2[k]));
		const float dmlp = net_to_density_d(raw[3], c.density_activation) * (dt * dot);
		dl[3] = ls * dmlp + (raw[3] < 0 ? -l1reg : 0.f);
		const __half2 lo = __floats2half2_rn(dl[0], dl[1]), hi = __floats2half2_rn(dl[2], dl[3]);
		uint2 pk; pk.x = *reinterpret_cast<const uint32_t*>(&lo); pk.y = *reinterpret_cast<const uint32_t*>(&hi);
		*reinterpret_cast<uint2*>(dloss + (size_t)(cbase + j) * padded_width) = pk;
	}
}

// fill_rollover_and_rescale (grads) + fill_rollover (coords) in one launch
__global__ void rollover_kernel(uint32_t batch, const uint32_t* __restrict__ n_in_ptr, __half* __restrict__ grads, uint32_t stride, Coordinate* __restrict__ coords) {
	const uint32_t n_in = *n_in_ptr;
	if (n_in == 0 || n_in >= batch) return;
	for (uint32_t i = n_in + blockIdx.x * blockDim.x + threadIdx.x; i < batch; i += gridDim.x * blockDim.x) {
		const uint32_t src = i % n_in;
		coords[i] = coords[src];
		for (uint32_t k = 0; k < stride; ++k)
			grads[(size_t)i * stride + k] = __float2half_rn(__half2float(grads[(size_t)src * stride + k]) * (float)n_in / (float)batch);
	}
}

} // namespace ngp

using namespace ngp;

extern "C" size_t ngp_composite_workspace_bytes(uint32_t n_rays) { return ngp_scan_workspace_bytes(n_rays); }

extern "C" int ngp_composite_train(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays, uint32_t max_compacted, const uint32_t* d_ray_counter,
                                   const float* d_rays_rgba, uint32_t* d_numsteps, const float* d_coords, float* d_coords_compacted,
                                   uint32_t* d_compacted_counter, const void* d_rgbsigma, void* d_dloss, uint32_t padded_width,
                                   const float* d_mean_density, uint64_t rng_state, uint64_t rng_inc, float* d_loss, void* d_scan_ws) {
	NGP_REQUIRE(cfg && d_ray_counter && d_rays_rgba && d_numsteps && d_coords && d_coords_compacted && d_compacted_counter && d_rgbsigma && d_dloss && d_mean_density && d_scan_ws, "null pointer");
	NGP_REQUIRE(padded_width >= 4 && padded_width % 4 == 0, "padded_width must be a multiple of 4");
	NGP_REQUIRE(n_rays > 0, "n_rays must be positive");
	cudaStream_t s = (cudaStream_t)stream;
	if (cudaMemsetAsync(d_scan_ws, 0, ngp_scan_workspace_bytes(n_rays), s) != cudaSuccess) { set_error("ngp_composite_train: memset failed"); return -2; }
	Pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
	ScanState scan{(unsigned long long*)d_scan_ws};
	NGP_LAUNCH("composite_train_kernel", s, composite_train_kernel<<<div_up(n_rays, 128), 128, 0, s>>>(*cfg, n_rays, max_compacted, d_ray_counter, (const float4*)d_rays_rgba, d_numsteps,
		(const Coordinate*)d_coords, (Coordinate*)d_coords_compacted, d_compacted_counter, (const __half*)d_rgbsigma, (__half*)d_dloss, padded_width,
		d_mean_density, rng, d_loss, scan));
	return check_launch("composite_train");
}

extern "C" int ngp_rollover(void* stream, uint32_t batch, const uint32_t* d_compacted_counter, void* d_dloss, uint32_t padded_width, float* d_coords_compacted) {
	NGP_REQUIRE(d_compacted_counter && d_dloss && d_coords_compacted, "null pointer");
	NGP_LAUNCH("rollover_kernel", (cudaStream_t)stream, rollover_kernel<<<grid_for(batch, 256), 256, 0, (cudaStream_t)stream>>>(batch, d_compacted_counter, (__half*)d_dloss, padded_width, (Coordinate*)d_coords_compacted));
	return check_launch("rollover");
}
