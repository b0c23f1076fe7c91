// Inference-side render kernels for sm_100a: per-iteration compositing with early termination, alive/hit compaction,
// final shade scatter, spp accumulation and tonemap.
// Replaces volume_rendering_kernel (internal/render/ray_marcher.cu:215-295), compact_rays_kernel and shade_kernel
// (internal/testbed.cu:423-477), accumulate_kernel / tonemap_kernel (internal/utils/render_buffer.cu:51-186; the
// cudaArray surface becomes a linear float4 device buffer -- headless, no graphics interop).
#include "composite.cuh"

namespace ngp {

// where the fused variant sends a ray after its compositing: alive -> the next ping-pong buffer, finished with alpha > 0.001 -> the hit
// list (compact_rays_kernel's rule, testbed.cu:423-446), so the next round needs no compaction launch of its own
struct CompactTargets { RayPayload* dst_p; float4* dst_c; RayPayload* hit_p; float4* hit_c; uint32_t* n_alive; uint32_t* n_hit; };

template <bool COMPACT>
__global__ void __launch_bounds__(128) composite_infer_kernel(
	ngp_render_cfg c, uint32_t n_rays_max, const uint32_t* __restrict__ n_rays_ptr, uint32_t n_steps_arg, uint32_t n_steps_offset_arg,
	const uint32_t* __restrict__ n_steps_ptr, const uint32_t* __restrict__ n_steps_offset_ptr,
	RayPayload* __restrict__ payloads, float4* __restrict__ rgba, const Coordinate* __restrict__ coords, const __half* __restrict__ net_out,
	uint32_t padded_width, int render_mode, float aabb_min, float aabb_max, float fwd_x, float fwd_y, float fwd_z, float depth_scale, CompactTargets ct)
{
	const uint32_t n_rays = n_rays_ptr ? min(*n_rays_ptr, n_rays_max) : n_rays_max;
	const uint32_t n_steps = n_steps_ptr ? *n_steps_ptr : n_steps_arg;
	const uint32_t n_steps_offset = n_steps_offset_ptr ? *n_steps_offset_ptr : n_steps_offset_arg;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (!COMPACT && i >= n_rays) return;
	const bool in_range = i < n_rays;
	RayPayload p;
	float4 l = make_float4(0.f, 0.f, 0.f, 0.f);
	if (in_range) { p = payloads[i]; l = rgba[i]; } else { p.alive = 0; p.n_steps = 0; }
	if (!COMPACT && !p.alive) return;
	const bool was_alive = in_range && p.alive;
	const uint32_t steps = was_alive ? p.n_steps : 0u;
	uint32_t j = 0;
	for (; j < steps; ++j) {
		const Coordinate co = coords[(size_t)i * n_steps + j];
		float raw[4];
		load_rgbsigma(net_out + ((size_t)i * n_steps + j) * padded_width, raw);
		const float dt = unwarp_dt(co.dt, c.min_step, c.n_cascades);
		const float T = 1.f - l.w;
		const float alpha = 1.f - __expf(-net_to_density(raw[3], c.density_activation) * dt);
		const float w = alpha * T;
		float r = net_to_rgb(raw[0], c.rgb_activation), g = net_to_rgb(raw[1], c.rgb_activation), b = net_to_rgb(raw[2], c.rgb_activation);
		if (render_mode != 1) {
			const float diag = aabb_max - aabb_min;
			const float px = aabb_min + co.px * diag, py = aabb_min + co.py * diag, pz = aabb_min + co.pz * diag;
			if (render_mode == 3) { const float z = (fwd_x * (px - p.ox) + fwd_y * (py - p.oy) + fwd_z * (pz - p.oz)) * depth_scale; r = g = b = z; }
			else if (render_mode == 4) { const float dx = px - p.ox, dy = py - p.oy, dz = pz - p.oz; const float z = sqrtf(dx * dx + dy * dy + dz * dz) * depth_scale; r = g = b = z; }
			else if (render_mode == 5) { r = g = b = co.dt; }
			else if (render_mode == 0) { r = g = b = alpha; }
			else if (render_mode == 2) {
				// ray_marcher.cu:260-266: the Coordinate holds d sigma_raw / d position (ngp_network_input_gradient ran on it); chain rule
				// through the density activation, the normal points towards decreasing density
				const float k = -net_to_density_d(raw[3], c.density_activation);
				const float nx = k * co.px, ny = k * co.py, nz = k * co.pz;
				// Eigen's normalized(): divide by the norm when the squared norm is positive, else leave the (zero) vector as it is
				const float n2 = nx * nx + ny * ny + nz * nz;
				r = nx; g = ny; b = nz;
				if (n2 > 0.f) { const float nn = sqrtf(n2); r = nx / nn; g = ny / nn; b = nz / nn; }
			}
		}
		l.x += r * w; l.y += g * w; l.z += b * w; l.w += w;
		if (l.w > (1.0f - c.transmittance_threshold)) break;
	}
	if (was_alive && j < n_steps) { p.alive = 0; p.n_steps = j + n_steps_offset; }
	if (!COMPACT) {
		payloads[i].alive = p.alive; payloads[i].n_steps = p.n_steps;
		rgba[i] = l;
		return;
	}
	// compaction, warp-aggregated: one atomicAdd per warp and list
	const uint32_t lane = threadIdx.x & 31;
	const bool alive = was_alive && p.alive != 0;
	const bool hit = in_range && !alive && l.w > 0.001f;
	const uint32_t am = __ballot_sync(0xffffffffu, alive), hm = __ballot_sync(0xffffffffu, hit);
	uint32_t abase = 0, hbase = 0;
	if (lane == 0) {
		if (am) abase = atomicAdd(ct.n_alive, __popc(am));
		if (hm) hbase = atomicAdd(ct.n_hit, __popc(hm));
	}
	abase = __shfl_sync(0xffffffffu, abase, 0); hbase = __shfl_sync(0xffffffffu, hbase, 0);
	const uint32_t below = (1u << lane) - 1u;
	if (alive) { const uint32_t d = abase + __popc(am & below); ct.dst_p[d] = p; ct.dst_c[d] = l; }
	else if (hit) { const uint32_t d = hbase + __popc(hm & below); ct.hit_p[d] = p; ct.hit_c[d] = l; }
}

// alive -> next ping-pong buffer, finished with alpha > 0.001 -> hit list.  Warp-aggregated: one atomicAdd per warp per
// list (ballot + popc prefix) instead of one per ray.
__global__ void __launch_bounds__(256) compact_rays_kernel(
	uint32_t n_rays_max, const uint32_t* __restrict__ n_rays_ptr, const RayPayload* __restrict__ src_p, const float4* __restrict__ src_c,
	RayPayload* __restrict__ dst_p, float4* __restrict__ dst_c, RayPayload* __restrict__ hit_p, float4* __restrict__ hit_c,
	uint32_t* __restrict__ n_alive, uint32_t* __restrict__ n_hit)
{
	const uint32_t n_rays = n_rays_ptr ? min(*n_rays_ptr, n_rays_max) : n_rays_max;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	const uint32_t lane = threadIdx.x & 31;
	RayPayload p; float4 col = make_float4(0.f, 0.f, 0.f, 0.f);
	bool alive = false, hit = false;
	if (i < n_rays) {
		p = src_p[i]; col = src_c[i];
		alive = p.alive != 0;
		hit = !alive && col.w > 0.001f;
	}
	const uint32_t am = __ballot_sync(0xffffffffu, alive), hm = __ballot_sync(0xffffffffu, hit);
	uint32_t abase = 0, hbase = 0;
	if (lane == 0) {
		if (am) abase = atomicAdd(n_alive, __popc(am));
		if (hm) hbase = atomicAdd(n_hit, __popc(hm));
	}
	abase = __shfl_sync(0xffffffffu, abase, 0); hbase = __shfl_sync(0xffffffffu, hbase, 0);
	const uint32_t below = (1u << lane) - 1u;
	if (alive) { const uint32_t d = abase + __popc(am & below); dst_p[d] = p; dst_c[d] = col; }
	else if (hit) { const uint32_t d = hbase + __popc(hm & below); hit_p[d] = p; hit_c[d] = col; }
}

// Round control of the device-driven render loop (the host loop of testbed.cu:521-575, one thread): after the compaction of round
// `round` it turns the live-ray count into the round's step count and sample count, keeps the running samples-per-ray offset and
// re-arms the counter the next compaction will fill.
//   ctl[0], ctl[1]: live rays, ping-pong (round r's compaction reads [r & 1] -- round 0 the pixel count -- and fills [(r + 1) & 1])
//   ctl[2] n_steps of this round, ctl[3] samples of this round, ctl[4] i_steps_offset (1 + samples per live ray so far)
__global__ void render_round_kernel(uint32_t* __restrict__ ctl, uint32_t round, uint32_t n_pixels, uint32_t min_steps, uint32_t max_steps, uint32_t tail_steps, uint32_t march_iter,
                                    uint32_t min_batch) {
	uint32_t offset = round == 0 ? 1u : ctl[4] + ctl[2];             // the previous round composited ctl[2] samples per live ray
	uint32_t n_alive = ctl[(round + 1) & 1];
	if (offset >= march_iter) n_alive = 0;                           // `while (n_alive > 0 && i_steps_offset < MARCH_ITER)`
	uint32_t n_steps = 0;
	if (n_alive) {
		// Our own extensions of the schedule, not of the result (a ray's samples and their compositing order do not depend on how
		// they are grouped into rounds).  (1) The reference aims at one sample per PIXEL and round, which starves small frames and
		// row bands (a 135-row band of a 1080p frame took 72 one-step rounds): the batch it aims at is at least min_batch samples.
		// (2) Once the batch has shrunk below a quarter of that, rounds are latency bound, so more steps are taken per round.
		const uint32_t batch = max(n_pixels, min_batch);
		n_steps = min(max(batch / n_alive, min_steps), max_steps);
		if (tail_steps > n_steps) n_steps = max(n_steps, min(tail_steps, batch / (4u * n_alive)));
	}
	ctl[(round + 1) & 1] = n_alive;
	ctl[round & 1] = 0;
	ctl[2] = n_steps; ctl[3] = n_alive * n_steps; ctl[4] = offset;
}

__global__ void shade_kernel(uint32_t n_hit_max, const uint32_t* __restrict__ n_hit_ptr, const float4* __restrict__ rgba, const RayPayload* __restrict__ payloads,
                             float4* __restrict__ frame, int render_mode, int train_in_linear) {
	const uint32_t n_hit = n_hit_ptr ? min(*n_hit_ptr, n_hit_max) : n_hit_max;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_hit) return;
	float4 t = rgba[i];
	const RayPayload p = payloads[i];
	if (render_mode == 2) {          // testbed.cu:463-465: accumulated normal -> unit length -> [0, 1], premultiplied
		const float n2 = t.x * t.x + t.y * t.y + t.z * t.z;      // Eigen's normalized(): a zero vector stays zero (no 0 / 0)
		if (n2 > 0.f) { const float nn = sqrtf(n2); t.x /= nn; t.y /= nn; t.z /= nn; }
		t.x = (0.5f * t.x + 0.5f) * t.w; t.y = (0.5f * t.y + 0.5f) * t.w; t.z = (0.5f * t.z + 0.5f) * t.w;
	}
	if (render_mode == 6) { const float col = (float)p.n_steps / 128; t = make_float4(col, col, col, 1.0f); }
	if (!train_in_linear && render_mode == 1) { t.x = srgb_to_linear(t.x); t.y = srgb_to_linear(t.y); t.z = srgb_to_linear(t.z); }
	const float4 f = frame[p.idx];
	frame[p.idx] = make_float4(t.x + f.x * (1.0f - t.w), t.y + f.y * (1.0f - t.w), t.z + f.z * (1.0f - t.w), t.w + f.w * (1.0f - t.w));
}

__global__ void accumulate_kernel(uint32_t n, const float4* __restrict__ frame, float4* __restrict__ accum, float sample_count, int color_space) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	float4 c = frame[i], t = accum[i];
	if (color_space == 1) { c.x = linear_to_srgb(c.x); c.y = linear_to_srgb(c.y); c.z = linear_to_srgb(c.z); }
	t.x = (t.x * sample_count + c.x) / (sample_count + 1); t.y = (t.y * sample_count + c.y) / (sample_count + 1);
	t.z = (t.z * sample_count + c.z) / (sample_count + 1); t.w = (t.w * sample_count + c.w) / (sample_count + 1);
	accum[i] = t;
}

__device__ inline void tonemap_curve(float* x, int curve) {
	if (curve == 0) return;
	for (int k = 0; k < 3; ++k) x[k] = fmaxf(x[k], 0.f);
	float k0, k1, k2, k3, k4, k5;
	if (curve == 1) {        // ACES approximation with pre-exposure folded in
		k0 = 0.6f * 0.6f * 2.51f; k1 = 0.6f * 0.03f; k2 = 0.0f; k3 = 0.6f * 0.6f * 2.43f; k4 = 0.6f * 0.59f; k5 = 0.14f;
	} else if (curve == 2) { // Hable filmic
		const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f;
		k0 = A * F - A * E; k1 = C * B * F - B * E; k2 = 0.0f; k3 = A * F; k4 = B * F; k5 = D * F * F;
		const float W = 11.2f, nom = k0 * (W * W) + k1 * W + k2, denom = k3 * (W * W) + k4 * W + k5, ws = denom / nom;
		k0 = 4.0f * k0 * ws; k1 = 2.0f * k1 * ws; k2 = k2 * ws; k3 = 4.0f * k3; k4 = 2.0f * k4;
	} else {                 // Reinhard on luminance
		const float Y = 0.2126f * x[0] + 0.7152f * x[1] + 0.0722f * x[2];
		for (int k = 0; k < 3; ++k) x[k] = x[k] * (1.f / (Y + 1.0f));
		return;
	}
	for (int k = 0; k < 3; ++k) { const float sq = x[k] * x[k]; x[k] = (sq * k0 + k1 * x[k] + k2) / (k3 * sq + k4 * x[k] + k5); }
}

__global__ void tonemap_kernel(uint32_t n, float exposure, float4 bg_in, const float4* __restrict__ accum, int color_space, int out_color_space, int curve, float4* __restrict__ out) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const float4 a = accum[i];
	const float scale = powf(2.0f, exposure);
	float c[3] = {a.x * scale, a.y * scale, a.z * scale};
	float bg[3] = {bg_in.x, bg_in.y, bg_in.z};
	if (color_space != 1) for (int k = 0; k < 3; ++k) bg[k] = srgb_to_linear(bg[k]);
	const float w = (1 - a.w) * bg_in.w;
	for (int k = 0; k < 3; ++k) c[k] += bg[k] * w;
	if (color_space == 1) for (int k = 0; k < 3; ++k) c[k] = srgb_to_linear(c[k]);
	tonemap_curve(c, curve);
	if (out_color_space == 1) for (int k = 0; k < 3; ++k) c[k] = linear_to_srgb(c[k]);
	out[i] = make_float4(c[0], c[1], c[2], a.w + w);
}

} // namespace ngp

using namespace ngp;

static int composite_infer_impl(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps, uint32_t n_steps_offset,
                                const uint32_t* d_n_steps, const uint32_t* d_n_steps_offset,
                                void* d_payloads, float* d_rgba, const float* d_coords, const void* d_rgbsigma, uint32_t padded_width, int render_mode,
                                float aabb_min, float aabb_max, const float* camera_fwd_host, float depth_scale, const CompactTargets* ct = nullptr) {
	NGP_REQUIRE(cfg && d_payloads && d_rgba && d_coords && d_rgbsigma && camera_fwd_host, "null pointer");
	if (n_rays == 0) return 0;
	cudaStream_t s = (cudaStream_t)stream;
	if (ct) {
		NGP_LAUNCH("composite_infer_kernel", s, composite_infer_kernel<true><<<div_up(n_rays, 128), 128, 0, s>>>(*cfg, n_rays, d_n_rays, n_steps, n_steps_offset,
			d_n_steps, d_n_steps_offset, (RayPayload*)d_payloads, (float4*)d_rgba, (const Coordinate*)d_coords, (const __half*)d_rgbsigma, padded_width, render_mode,
			aabb_min, aabb_max, camera_fwd_host[0], camera_fwd_host[1], camera_fwd_host[2], depth_scale, *ct));
	} else {
		NGP_LAUNCH("composite_infer_kernel", s, composite_infer_kernel<false><<<div_up(n_rays, 128), 128, 0, s>>>(*cfg, n_rays, d_n_rays, n_steps, n_steps_offset,
			d_n_steps, d_n_steps_offset, (RayPayload*)d_payloads, (float4*)d_rgba, (const Coordinate*)d_coords, (const __half*)d_rgbsigma, padded_width, render_mode,
			aabb_min, aabb_max, camera_fwd_host[0], camera_fwd_host[1], camera_fwd_host[2], depth_scale, CompactTargets{}));
	}
	return check_launch("composite_infer");
}

extern "C" int ngp_composite_infer(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps, uint32_t n_steps_offset,
                                   void* d_payloads, float* d_rgba, const float* d_coords, const void* d_rgbsigma, uint32_t padded_width, int render_mode,
                                   float aabb_min, float aabb_max, const float* camera_fwd_host, float depth_scale) {
	return composite_infer_impl(stream, cfg, n_rays, d_n_rays, n_steps, n_steps_offset, nullptr, nullptr, d_payloads, d_rgba, d_coords, d_rgbsigma, padded_width, render_mode,
	                            aabb_min, aabb_max, camera_fwd_host, depth_scale);
}

extern "C" int ngp_composite_infer_dev(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays_bound, const uint32_t* d_ctl, uint32_t round,
                                       void* d_payloads, float* d_rgba, const float* d_coords, const void* d_rgbsigma, uint32_t padded_width, int render_mode,
                                       float aabb_min, float aabb_max, const float* camera_fwd_host, float depth_scale) {
	NGP_REQUIRE(d_ctl, "null pointer");
	return composite_infer_impl(stream, cfg, n_rays_bound, d_ctl + ((round + 1) & 1), 0, 0, d_ctl + 2, d_ctl + 4, d_payloads, d_rgba, d_coords, d_rgbsigma, padded_width,
	                            render_mode, aabb_min, aabb_max, camera_fwd_host, depth_scale);
}

extern "C" int ngp_composite_infer_compact_dev(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays_bound, uint32_t* d_ctl, uint32_t round,
                                               void* d_payloads, float* d_rgba, const float* d_coords, const void* d_rgbsigma, uint32_t padded_width, int render_mode,
                                               float aabb_min, float aabb_max, const float* camera_fwd_host, float depth_scale,
                                               void* d_next_payloads, float* d_next_rgba, void* d_hit_payloads, float* d_hit_rgba, uint32_t* d_n_hit) {
	NGP_REQUIRE(d_ctl && d_next_payloads && d_next_rgba && d_hit_payloads && d_hit_rgba && d_n_hit, "null pointer");
	// the live rays of round + 1 are counted where that round's compaction would have counted them: d_ctl[(round + 2) & 1],
	// which ngp_render_round(round) has cleared
	const CompactTargets ct{(RayPayload*)d_next_payloads, (float4*)d_next_rgba, (RayPayload*)d_hit_payloads, (float4*)d_hit_rgba, d_ctl + (round & 1), d_n_hit};
	return composite_infer_impl(stream, cfg, n_rays_bound, d_ctl + ((round + 1) & 1), 0, 0, d_ctl + 2, d_ctl + 4, d_payloads, d_rgba, d_coords, d_rgbsigma, padded_width,
	                            render_mode, aabb_min, aabb_max, camera_fwd_host, depth_scale, &ct);
}

extern "C" int ngp_render_round(void* stream, uint32_t* d_ctl, uint32_t round, uint32_t n_pixels, uint32_t min_steps, uint32_t max_steps, uint32_t tail_steps, uint32_t march_iter,
                                uint32_t min_batch) {
	NGP_REQUIRE(d_ctl, "null pointer");
	NGP_REQUIRE(min_steps >= 1 && max_steps >= min_steps, "bad step limits");
	NGP_LAUNCH("render_round_kernel", (cudaStream_t)stream, render_round_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(d_ctl, round, n_pixels, min_steps, max_steps, tail_steps, march_iter, min_batch));
	return check_launch("render_round");
}

extern "C" int ngp_compact_rays(void* stream, uint32_t n_rays, const uint32_t* d_n_rays, const void* d_src_payloads, const float* d_src_rgba,
                                void* d_dst_payloads, float* d_dst_rgba, void* d_hit_payloads, float* d_hit_rgba, uint32_t* d_n_alive, uint32_t* d_n_hit) {
	NGP_REQUIRE(d_src_payloads && d_src_rgba && d_dst_payloads && d_dst_rgba && d_hit_payloads && d_hit_rgba && d_n_alive && d_n_hit, "null pointer");
	if (n_rays == 0) return 0;
	NGP_LAUNCH("compact_rays_kernel", (cudaStream_t)stream, compact_rays_kernel<<<div_up(n_rays, 256), 256, 0, (cudaStream_t)stream>>>(n_rays, d_n_rays, (const RayPayload*)d_src_payloads, (const float4*)d_src_rgba,
		(RayPayload*)d_dst_payloads, (float4*)d_dst_rgba, (RayPayload*)d_hit_payloads, (float4*)d_hit_rgba, d_n_alive, d_n_hit));
	return check_launch("compact_rays");
}

extern "C" int ngp_shade(void* stream, uint32_t n_hit_max, const uint32_t* d_n_hit, const float* d_rgba, const void* d_payloads, float* d_frame, int render_mode, int train_in_linear) {
	NGP_REQUIRE(d_rgba && d_payloads && d_frame, "null pointer");
	if (n_hit_max == 0) return 0;
	NGP_LAUNCH("shade_kernel", (cudaStream_t)stream, shade_kernel<<<div_up(n_hit_max, 256), 256, 0, (cudaStream_t)stream>>>(n_hit_max, d_n_hit, (const float4*)d_rgba, (const RayPayload*)d_payloads, (float4*)d_frame, render_mode, train_in_linear));
	return check_launch("shade");
}

extern "C" int ngp_accumulate(void* stream, uint32_t n_pixels, const float* d_frame, float* d_accum, float sample_count, int color_space) {
	NGP_REQUIRE(d_frame && d_accum, "null pointer");
	if (n_pixels == 0) return 0;
	NGP_LAUNCH("accumulate_kernel", (cudaStream_t)stream, accumulate_kernel<<<div_up(n_pixels, 256), 256, 0, (cudaStream_t)stream>>>(n_pixels, (const float4*)d_frame, (float4*)d_accum, sample_count, color_space));
	return check_launch("accumulate");
}

extern "C" int ngp_tonemap(void* stream, uint32_t n_pixels, float exposure, const float* background_rgba_host, const float* d_accum, int color_space,
                           int output_color_space, int tonemap_curve_id, float* d_out) {
	NGP_REQUIRE(background_rgba_host && d_accum && d_out, "null pointer");
	if (n_pixels == 0) return 0;
	const float4 bg = make_float4(background_rgba_host[0], background_rgba_host[1], background_rgba_host[2], background_rgba_host[3]);
	NGP_LAUNCH("tonemap_kernel", (cudaStream_t)stream, tonemap_kernel<<<div_up(n_pixels, 256), 256, 0, (cudaStream_t)stream>>>(n_pixels, exposure, bg, (const float4*)d_accum, color_space, output_color_space, tonemap_curve_id, (float4*)d_out));
	return check_launch("tonemap");
}
