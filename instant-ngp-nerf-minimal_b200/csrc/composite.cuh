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
