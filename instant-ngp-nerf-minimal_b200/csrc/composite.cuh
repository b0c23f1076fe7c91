// Activation, loss and step-size helpers shared by the compositing kernels.
// internal/utils/common_device.h:292-342 (activations), :395-404 (dt warp), internal/utils/loss.cu:23-103 (losses).
#pragma once
#include "common.cuh"

namespace ngp {

enum { ACT_NONE = 0, ACT_RELU = 1, ACT_LOGISTIC = 2, ACT_EXP = 3 };
enum { LOSS_L2 = 0, LOSS_L1, LOSS_MAPE, LOSS_SMAPE, LOSS_SMOOTHL1, LOSS_LOGL1, LOSS_RELL2 };

__device__ __forceinline__ float clamp_f(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
__device__ __forceinline__ float logistic(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ float net_to_rgb(float v, int act) {
	switch (act) {
		case ACT_NONE: return v;
		case ACT_RELU: return v > 0.f ? v : 0.f;
		case ACT_LOGISTIC: return logistic(v);
		default: return __expf(clamp_f(v, -10.f, 10.f));
	}
}
__device__ __forceinline__ float net_to_rgb_d(float v, int act) {
	switch (act) {
		case ACT_NONE: return 1.f;
		case ACT_RELU: return v > 0.f ? 1.f : 0.f;
		case ACT_LOGISTIC: { const float s = logistic(v); return s * (1 - s); }
		default: return __expf(clamp_f(v, -10.f, 10.f));
	}
}
__device__ __forceinline__ float net_to_density(float v, int act) {
	switch (act) {
		case ACT_NONE: return v;
		case ACT_RELU: return v > 0.f ? v : 0.f;
		case ACT_LOGISTIC: return logistic(v);
		default: return __expf(v);
	}
}
__device__ __forceinline__ float net_to_density_d(float v, int act) {
	switch (act) {
		case ACT_NONE: return 1.f;
		case ACT_RELU: return v > 0.f ? 1.f : 0.f;
		case ACT_LOGISTIC: { const float s = logistic(v); return s * (1 - s); }
		default: return __expf(clamp_f(v, -15.f, 15.f));
	}
}

__device__ __forceinline__ void loss_and_grad(int type, const float* tgt, const float* pred, float* loss, float* grad) {
	#pragma unroll
	for (int k = 0; k < 3; ++k) {
		const float d = pred[k] - tgt[k];
		switch (type) {
			case LOSS_RELL2: { const float f = 1.f / (pred[k] * pred[k] + 1e-2f); loss[k] = d * d * f; grad[k] = 2.f * d * f; break; }
			case LOSS_L1: loss[k] = fabsf(d); grad[k] = copysignf(1.f, d); break;
			case LOSS_MAPE: { const float f = 1.f / (fabsf(pred[k]) + 1e-2f); loss[k] = fabsf(d) * f; grad[k] = copysignf(f, d); break; }
			case LOSS_SMAPE: { const float f = 1.f / (0.5f * (fabsf(pred[k]) + fabsf(tgt[k])) + 1e-2f); loss[k] = fabsf(d) * f; grad[k] = copysignf(f, d); break; }
			case LOSS_SMOOTHL1: {
				const float a = 0.1f, ad = fabsf(d);
				loss[k] = ad > a ? (ad - 0.5f * a) : (0.5f / a * d * d);
				grad[k] = ad > a ? (d > 0 ? 1.0f : -1.0f) : (d / a);
				break;
			}
			case LOSS_LOGL1: { const float dv = fabsf(d) + 1.f; loss[k] = logf(dv); grad[k] = copysignf(1.f / dv, d); break; }
			default: loss[k] = d * d; grad[k] = 2.f * d; break;
		}
	}
}

__device__ __forceinline__ float unwarp_dt(float dt, float min_step, uint32_t n_cascades) {
	const float max_stepsize = min_step * (float)(1u << (n_cascades - 1));
	return dt * (max_stepsize - min_step) + min_step;
}

// ---- trainable environment map (internal/utils/envmap.h:19-88) -----------------------------------------------------------
// texel coordinates and bilinear weights of a direction: dir_to_spherical_unorm({d.z, -d.x, d.y}) (utils/random_val.cuh:47-52),
// x from the azimuth, y from the polar angle; the x index wraps, the y index clamps
struct EnvmapTap { int x0, y0; float wx, wy; };
__device__ __forceinline__ EnvmapTap envmap_tap(float dx, float dy, float dz, int res_x, int res_y) {
	const float vx = dz, vy = -dx, vz = dy;
	const float cos_theta = fminf(fmaxf(vz, -1.0f), 1.0f);
	const float theta = acosf(cos_theta);
	const float phi = atan2f(vy, vx);
	const float cx = theta / 3.14159265358979323846f, cy = phi / (2.0f * 3.14159265358979323846f) + 0.5f;
	const float fx = cy * (float)(res_x - 1), fy = cx * (float)(res_y - 1);
	EnvmapTap t;
	t.x0 = (int)fx; t.y0 = (int)fy;
	t.wx = fx - (float)t.x0; t.wy = fy - (float)t.y0;
	return t;
}
__device__ __forceinline__ size_t envmap_texel(int x, int y, int res_x, int res_y) {
	if (x < 0) x += res_x; else if (x >= res_x) x -= res_x;
	y = max(min(y, res_y - 1), 0);
	return ((size_t)x + (size_t)y * res_x) * 4;
}
__device__ __forceinline__ void read_envmap(const __half* __restrict__ env, int res_x, int res_y, float dx, float dy, float dz, float* out) {
	const EnvmapTap t = envmap_tap(dx, dy, dz, res_x, res_y);
	const float w[4] = {(1 - t.wx) * (1 - t.wy), t.wx * (1 - t.wy), (1 - t.wx) * t.wy, t.wx * t.wy};
	out[0] = out[1] = out[2] = out[3] = 0.f;
	#pragma unroll
	for (int k = 0; k < 4; ++k) {
		const uint2 v = *reinterpret_cast<const uint2*>(env + envmap_texel(t.x0 + (k & 1), t.y0 + (k >> 1), res_x, res_y));
		float tex[4];
		{ const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&v.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&v.y)); tex[0] = a.x; tex[1] = a.y; tex[2] = b.x; tex[3] = b.y; }
		#pragma unroll
		for (int c = 0; c < 4; ++c) out[c] += w[k] * tex[c];
	}
}
// value and weight are rounded to fp16 and multiplied in fp16 like deposit_val (envmap.h:63-80), two half2 atomics per texel
__device__ __forceinline__ void deposit_envmap_gradient(const float* value, __half* __restrict__ grad, int res_x, int res_y, float dx, float dy, float dz) {
	const EnvmapTap t = envmap_tap(dx, dy, dz, res_x, res_y);
	const float w[4] = {(1 - t.wx) * (1 - t.wy), t.wx * (1 - t.wy), (1 - t.wx) * t.wy, t.wx * t.wy};
	const __half2 v01 = __floats2half2_rn(value[0], value[1]), v23 = __floats2half2_rn(value[2], value[3]);
	#pragma unroll
	for (int k = 0; k < 4; ++k) {
		const __half2 wk = __float2half2_rn(w[k]);
		__half* p = grad + envmap_texel(t.x0 + (k & 1), t.y0 + (k >> 1), res_x, res_y);
		atomicAdd(reinterpret_cast<__half2*>(p), __hmul2(v01, wk));
		atomicAdd(reinterpret_cast<__half2*>(p + 2), __hmul2(v23, wk));
	}
}
__device__ __forceinline__ float srgb_to_linear_derivative(float srgb) {
	return srgb <= 0.04045f ? 1.0f / 12.92f : 2.4f / 1.055f * powf((srgb + 0.055f) / 1.055f, 1.4f);
}

__device__ __forceinline__ void unpack_rgbsigma(const uint2 v, float* raw) {
	const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&v.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
	raw[0] = a.x; raw[1] = a.y; raw[2] = b.x; raw[3] = b.y;
}

__device__ __forceinline__ void load_rgbsigma(const __half* p, float* raw) {
	const uint2 v = *reinterpret_cast<const uint2*>(p);
	const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&v.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
	raw[0] = a.x; raw[1] = a.y; raw[2] = b.x; raw[3] = b.y;
}

} // namespace ngp
