// TEST INFRASTRUCTURE ONLY (see oracle_common.h).  Pinned to golden vectors from the reference's CUDA build (tests/golden/).
// CPU restatement of volume-render compositing:
//   training fwd + loss + bwd + compaction : internal/render/ray_marcher.cu:15-213
//   losses                                 : internal/utils/loss.cu:23-103
//   activations                            : internal/utils/common_device.h:292-342
//   rollover                               : tiny-cuda-nn/common_device.h:485-505
//   inference compositing                  : ray_marcher.cu:215-295
//   shade / accumulate / tonemap           : internal/testbed.cu:448-477, utils/render_buffer.cu:51-186
// expf stands in for the device __expf: float results carry a tolerance, counters are exact.
#include "oracle_common.h"
#include <omp.h>

using namespace orc;

namespace {

enum { ACT_NONE = 0, ACT_RELU = 1, ACT_LOGISTIC = 2, ACT_EXP = 3 };
enum { LOSS_L2 = 0, LOSS_L1, LOSS_MAPE, LOSS_SMAPE, LOSS_SMOOTHL1, LOSS_LOGL1, LOSS_RELL2 };

static inline float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
static inline float logistic(float x) { return 1.0f / (1.0f + expf(-x)); }

static inline float net_to_rgb(float v, int act) {
	switch (act) {
		case ACT_NONE: return v;
		case ACT_RELU: return v > 0.f ? v : 0.f;
		case ACT_LOGISTIC: return logistic(v);
		default: return expf(clampf(v, -10.f, 10.f));
	}
}
static inline float net_to_rgb_d(float v, int act) {
	switch (act) {
		case ACT_NONE: return 1.f;
		case ACT_RELU: return v > 0.f ? 1.f : 0.f;
		case ACT_LOGISTIC: { float s = logistic(v); return s * (1 - s); }
		default: return expf(clampf(v, -10.f, 10.f));
	}
}
static inline float net_to_density(float v, int act) {
	switch (act) {
		case ACT_NONE: return v;
		case ACT_RELU: return v > 0.f ? v : 0.f;
		case ACT_LOGISTIC: return logistic(v);
		default: return expf(v);
	}
}
static inline float net_to_density_d(float v, int act) {
	switch (act) {
		case ACT_NONE: return 1.f;
		case ACT_RELU: return v > 0.f ? 1.f : 0.f;
		case ACT_LOGISTIC: { float s = logistic(v); return s * (1 - s); }
		default: return expf(clampf(v, -15.f, 15.f));
	}
}

static inline void loss_and_grad(int type, const float* tgt, const float* pred, float* loss, float* grad) {
	for (int k = 0; k < 3; ++k) {
		float d = pred[k] - tgt[k];
		switch (type) {
			case LOSS_RELL2: { float f = 1.f / (pred[k] * pred[k] + 1e-2f); loss[k] = d * d * f; grad[k] = 2.f * d * f; break; }
			case LOSS_L1: loss[k] = fabsf(d); grad[k] = copysignf(1.f, d); break;
			case LOSS_MAPE: { float f = 1.f / (fabsf(pred[k]) + 1e-2f); loss[k] = fabsf(d) * f; grad[k] = copysignf(f, d); break; }
			case LOSS_SMAPE: { float f = 1.f / (0.5f * (fabsf(pred[k]) + fabsf(tgt[k])) + 1e-2f); loss[k] = fabsf(d) * f; grad[k] = copysignf(f, d); break; }
			case LOSS_SMOOTHL1: {
				const float a = 0.1f; float ad = fabsf(d);
				loss[k] = ad > a ? (ad - 0.5f * a) : (0.5f / a * d * d);
				grad[k] = ad > a ? (d > 0 ? 1.0f : -1.0f) : (d / a);
				break;
			}
			case LOSS_LOGL1: { float dv = fabsf(d) + 1.f; loss[k] = logf(dv); grad[k] = copysignf(1.f / dv, d); break; }
			default: loss[k] = d * d; grad[k] = 2.f * d; break;
		}
	}
}

static inline float unwarp_dt(float dt, float min_step, uint32_t n_cascades) {
	float max_stepsize = min_step * (float)(1u << (n_cascades - 1));
	return dt * (max_stepsize - min_step) + min_step;
}

} // namespace

extern "C" {

// cfg_f: min_step, mean_density, min_optical_thickness, transmittance_threshold, loss_scale, bg[3]
// cfg_u: n_cascades, rgb_act, density_act, loss_type, color_space(0 linear,1 srgb), random_bg, train_in_linear
// numsteps [n_rays_kept][2] in/out.  Compaction in ray order.  Returns the final compacted counter
// (may exceed max_compacted, as in the reference).
// ---- environment map (internal/utils/envmap.h:19-88; dir_to_spherical_unorm: internal/utils/random_val.cuh:47-52) ----------
struct EnvTap { int x0, y0; float wx, wy; };
static EnvTap env_tap(const float* d, int res_x, int res_y) {
	const float v[3] = {d[2], -d[0], d[1]};                                   // envmap.h:20
	const float cos_theta = fminf(fmaxf(v[2], -1.0f), 1.0f);
	const float theta = acosf(cos_theta), phi = atan2f(v[1], v[0]);
	const float cx = theta / 3.14159265358979323846f, cy = phi / (2.0f * 3.14159265358979323846f) + 0.5f;
	const float fx = cy * (float)(res_x - 1), fy = cx * (float)(res_y - 1);   // envmap.h:22 (x from the azimuth)
	EnvTap t; t.x0 = (int)fx; t.y0 = (int)fy; t.wx = fx - (float)t.x0; t.wy = fy - (float)t.y0;
	return t;
}
static size_t env_texel(int x, int y, int res_x, int res_y) {                   // envmap.h:27-33
	if (x < 0) x += res_x; else if (x >= res_x) x -= res_x;
	y = std::max(std::min(y, res_y - 1), 0);
	return ((size_t)x + (size_t)y * res_x) * 4;
}
static void env_read(const uint16_t* env, int res_x, int res_y, const float* d, float* out) {
	const EnvTap t = env_tap(d, res_x, res_y);
	const float w[4] = {(1 - t.wx) * (1 - t.wy), t.wx * (1 - t.wy), (1 - t.wx) * t.wy, t.wx * t.wy};
	for (int c = 0; c < 4; ++c) out[c] = 0.f;
	for (int k = 0; k < 4; ++k) {
		const uint16_t* p = env + env_texel(t.x0 + (k & 1), t.y0 + (k >> 1), res_x, res_y);
		for (int c = 0; c < 4; ++c) out[c] += w[k] * h16from(p[c]);
	}
}
// the reference deposits fp16 products with fp16 atomics (order-dependent); the oracle sums the same fp16-rounded products in fp32
static void env_deposit(const float* value, float* grad, int res_x, int res_y, const float* d) {
	const EnvTap t = env_tap(d, res_x, res_y);
	const float w[4] = {(1 - t.wx) * (1 - t.wy), t.wx * (1 - t.wy), (1 - t.wx) * t.wy, t.wx * t.wy};
	for (int k = 0; k < 4; ++k) {
		float* p = grad + env_texel(t.x0 + (k & 1), t.y0 + (k >> 1), res_x, res_y);
		const float wk = h16from(h16bits(w[k]));
		for (int c = 0; c < 4; ++c) p[c] += h16from(h16bits(h16from(h16bits(value[c])) * wk));
	}
}
static float srgb_to_linear_derivative(float srgb) { return srgb <= 0.04045f ? 1.0f / 12.92f : 2.4f / 1.055f * powf((srgb + 0.055f) / 1.055f, 1.4f); }

static uint32_t composite_train_impl(uint32_t n_rays, uint32_t n_rays_kept, uint32_t max_compacted, const float* rgba_gt,
                             uint32_t* numsteps, const float* coords, float* coords_compacted,
                             const uint16_t* net_out, uint16_t* dloss, uint32_t padded_width,
                             const float* cfg_f, const uint32_t* cfg_u, uint64_t rng_state, uint64_t rng_inc,
                             float* loss_out, float* rgb_ray_out,
                             const float* rays, const uint16_t* envmap, float* envmap_grad, int env_res_x, int env_res_y, int env_loss_type) {
	const float min_step = cfg_f[0], mean_density = cfg_f[1], min_thick = cfg_f[2], T_thresh = cfg_f[3], loss_scale0 = cfg_f[4];
	const uint32_t n_cascades = cfg_u[0]; const int rgb_act = cfg_u[1], den_act = cfg_u[2], loss_type = cfg_u[3];
	const int color_space = cfg_u[4], random_bg = cfg_u[5], lin = cfg_u[6];
	float bg0[3] = {cfg_f[5], cfg_f[6], cfg_f[7]};
	if (random_bg) { Pcg32 r; r.state = rng_state; r.inc = rng_inc; bg0[0] = r.next_float(); bg0[1] = r.next_float(); bg0[2] = r.next_float(); }
	for (int k = 0; k < 3; ++k) bg0[k] = srgb_to_linear(bg0[k]);
	uint32_t counter = 0;
	for (uint32_t i = 0; i < n_rays_kept; ++i) {
		const uint32_t numst = numsteps[i * 2], base = numsteps[i * 2 + 1];
		const Coordinate* cs = (const Coordinate*)coords + base;
		const uint16_t* no = net_out + (size_t)base * 4;
		float T = 1.f, rgb_ray[3] = {0, 0, 0};
		uint32_t n_used = 0;
		for (; n_used < numst; ++n_used) {
			if (T < T_thresh) break;
			const uint16_t* o = no + (size_t)n_used * 4;
			float density = net_to_density(h16from(o[3]), den_act);
			float dt = unwarp_dt(cs[n_used].dt, min_step, n_cascades);
			float alpha = 1.f - expf(-density * dt), w = alpha * T;
			for (int k = 0; k < 3; ++k) rgb_ray[k] += w * net_to_rgb(h16from(o[k]), rgb_act);
			T *= (1.f - alpha);
		}
		float bg[3] = {bg0[0], bg0[1], bg0[2]}, tgt[3];
		if (envmap) {                                                           // ray_marcher.cu:96-102
			float ev[4];
			env_read(envmap, env_res_x, env_res_y, rays + (size_t)i * 6 + 3, ev);
			for (int k = 0; k < 3; ++k) bg[k] = ev[k] + bg[k] * (1.0f - ev[3]);
		}
		const float* gt = rgba_gt + (size_t)i * 4;
		if (lin || color_space == 0) {
			for (int k = 0; k < 3; ++k) tgt[k] = gt[k] + (1 - gt[3]) * bg[k];
			if (!lin) for (int k = 0; k < 3; ++k) { tgt[k] = linear_to_srgb(tgt[k]); bg[k] = linear_to_srgb(bg[k]); }
		} else {
			for (int k = 0; k < 3; ++k) bg[k] = linear_to_srgb(bg[k]);
			for (int k = 0; k < 3; ++k) tgt[k] = gt[3] > 0 ? linear_to_srgb(gt[k] / gt[3]) * gt[3] + (1 - gt[3]) * bg[k] : bg[k];
		}
		if (n_used == numst) for (int k = 0; k < 3; ++k) rgb_ray[k] += T * bg[k];
		if (rgb_ray_out) for (int k = 0; k < 3; ++k) rgb_ray_out[(size_t)i * 3 + k] = rgb_ray[k];
		uint32_t cbase = counter; counter += n_used;
		uint32_t n = std::min(max_compacted - std::min(max_compacted, cbase), n_used);
		numsteps[i * 2] = n; numsteps[i * 2 + 1] = cbase;
		if (n == 0) continue;
		float loss[3], g[3];
		loss_and_grad(loss_type, tgt, rgb_ray, loss, g);
		if (loss_out) loss_out[i] = ((loss[0] + loss[1] + loss[2]) / 3.f) / (float)n_rays;
		const float ls = loss_scale0 / (float)n_rays;
		const float l2reg = rgb_act == ACT_EXP ? 1e-4f : 0.f;
		const float l1reg = mean_density < min_thick ? 1e-4f : 0.f;
		float rgb2[3] = {0, 0, 0}; T = 1.f;
		for (uint32_t j = 0; j < n; ++j) {
			((Coordinate*)coords_compacted)[cbase + j] = cs[j];
			const uint16_t* o = no + (size_t)j * 4;
			float dt = unwarp_dt(cs[j].dt, min_step, n_cascades);
			float raw[4] = {h16from(o[0]), h16from(o[1]), h16from(o[2]), h16from(o[3])};
			float rgb[3] = {net_to_rgb(raw[0], rgb_act), net_to_rgb(raw[1], rgb_act), net_to_rgb(raw[2], rgb_act)};
			float density = net_to_density(raw[3], den_act);
			float alpha = 1.f - expf(-density * dt), w = alpha * T;
			for (int k = 0; k < 3; ++k) rgb2[k] += w * rgb[k];
			T *= (1.f - alpha);
			uint16_t* dl = dloss + (size_t)(cbase + j) * padded_width;
			for (int k = 0; k < 3; ++k)
				dl[k] = h16bits(ls * (w * g[k] * net_to_rgb_d(raw[k], rgb_act) + fmaxf(0.f, l2reg * raw[k])));
			float dot = 0.f;
			for (int k = 0; k < 3; ++k) dot += g[k] * (T * rgb[k] - (rgb_ray[k] - rgb2[k]));
			float dmlp = net_to_density_d(raw[3], den_act) * (dt * dot);
			dl[3] = h16bits(ls * dmlp + (raw[3] < 0 ? -l1reg : 0.f));
		}
		if (envmap_grad && n == numst) {                                        // ray_marcher.cu:190-212 (T = transmittance behind the ray)
			float ge[3] = {g[0], g[1], g[2]}, l2[3];
			if (env_loss_type != loss_type) loss_and_grad(env_loss_type, tgt, rgb_ray, l2, ge);
			float v[4];
			for (int k = 0; k < 3; ++k) {
				float d = T * ge[k];
				if (!lin) d /= srgb_to_linear_derivative(bg[k]);
				v[k] = ls * d;
			}
			v[3] = 0.f;
			env_deposit(v, envmap_grad, env_res_x, env_res_y, rays + (size_t)i * 6 + 3);
		}
	}
	return counter;
}

uint32_t orc_composite_train(uint32_t n_rays, uint32_t n_rays_kept, uint32_t max_compacted, const float* rgba_gt,
                             uint32_t* numsteps, const float* coords, float* coords_compacted,
                             const uint16_t* net_out, uint16_t* dloss, uint32_t padded_width,
                             const float* cfg_f, const uint32_t* cfg_u, uint64_t rng_state, uint64_t rng_inc,
                             float* loss_out, float* rgb_ray_out) {
	return composite_train_impl(n_rays, n_rays_kept, max_compacted, rgba_gt, numsteps, coords, coords_compacted, net_out, dloss, padded_width, cfg_f, cfg_u,
	                            rng_state, rng_inc, loss_out, rgb_ray_out, nullptr, nullptr, nullptr, 0, 0, 0);
}

// the same with a trainable environment map behind the volume: rays = 6 floats per kept ray {o, d}, envmap fp16 [res_y][res_x][4],
// envmap_grad fp32 of the same shape (accumulated, caller zeroes)
uint32_t orc_composite_train_envmap(uint32_t n_rays, uint32_t n_rays_kept, uint32_t max_compacted, const float* rgba_gt,
                             uint32_t* numsteps, const float* coords, float* coords_compacted,
                             const uint16_t* net_out, uint16_t* dloss, uint32_t padded_width,
                             const float* cfg_f, const uint32_t* cfg_u, uint64_t rng_state, uint64_t rng_inc,
                             float* loss_out, float* rgb_ray_out,
                             const float* rays, const uint16_t* envmap, float* envmap_grad, int32_t env_res_x, int32_t env_res_y, int32_t env_loss_type) {
	return composite_train_impl(n_rays, n_rays_kept, max_compacted, rgba_gt, numsteps, coords, coords_compacted, net_out, dloss, padded_width, cfg_f, cfg_u,
	                            rng_state, rng_inc, loss_out, rgb_ray_out, rays, envmap, envmap_grad, env_res_x, env_res_y, env_loss_type);
}

// fill_rollover(_and_rescale): grads [B][stride] fp16, coords [B][7] f32
void orc_rollover(uint32_t B, uint32_t n_in, uint16_t* grads, uint32_t stride, float* coords) {
	if (n_in == 0) return;
	for (uint64_t i = (uint64_t)n_in * stride; i < (uint64_t)B * stride; ++i)
		grads[i] = h16bits(h16from(grads[i % ((uint64_t)n_in * stride)]) * (float)n_in / (float)B);
	for (uint32_t i = n_in; i < B; ++i) std::memcpy(coords + (size_t)i * 7, coords + (size_t)(i % n_in) * 7, 28);
}

} // extern "C"

#include "render_infer.inl"
