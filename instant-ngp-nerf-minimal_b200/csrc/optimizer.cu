// Parameter initialisation and the fused Adam + EMA step for sm_100a.
// Replaces Trainer::initialize_params (tiny-cuda-nn/trainer.h:73-112), generate_random_kernel (random.h:69-94),
// adam_step (optimizers/adam.h:50-113) and ema_step_half_precision (optimizers/ema.h:65-79): one pass over the arena
// instead of two, 16-byte vector accesses, and the EMA of an untouched grid entry is still updated (it decays towards the
// unchanged weight exactly like the reference).  HBM-bound: 42 B / parameter when touched, 6 B when skipped.
#include "common.cuh"
#include <vector>
#include <cmath>

namespace ngp {

// reference pattern: thread i advances the generator by 4*i and writes elements i + total_threads*j, j < 4
__global__ void init_grid_kernel(uint64_t n, uint64_t n_threads, uint64_t total_threads, Pcg32 rng0, float* __restrict__ w32, __half* __restrict__ w16) {
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n_threads; i += (uint64_t)gridDim.x * blockDim.x) {
		Pcg32 rng = rng0;
		rng.advance(i * 4);
		for (uint64_t j = 0; j < 4; ++j) {
			const uint64_t idx = i + total_threads * j;
			if (idx >= n) break;
			const float v = rng.next_float() * (1e-4f - (-1e-4f)) + (-1e-4f);
			w32[idx] = v;
			w16[idx] = __float2half_rn(v);
		}
	}
}

struct AdamHyper { float lr, beta1, beta2, eps, l2_reg, loss_scale, ema_decay, deb_old, deb_new; };

__device__ __forceinline__ void adam_one(const AdamHyper& h, bool matrix, float gh, float& w, float& m1, float& m2, uint32_t& steps, __half& w16, bool& touched) {
	float g = gh / h.loss_scale;
	touched = matrix || g != 0.f;
	if (!touched) return;
	if (matrix) g += h.l2_reg * w;
	const float gsq = g * g;
	m1 = h.beta1 * m1 + (1 - h.beta1) * g;
	m2 = h.beta2 * m2 + (1 - h.beta2) * gsq;
	const uint32_t cs = ++steps;
	const float lr = h.lr * (sqrtf(1 - powf(h.beta2, (float)cs)) / (1 - powf(h.beta1, (float)cs)));
	const float eff = fminf(fmaxf(lr / (sqrtf(m2) + h.eps), 0.f), 3.402823466e38f);
	w = w - eff * m1;
	w16 = __float2half_rn(w);
}

// 4 parameters per thread (n is padded up by the caller's allocation to a multiple of 4 only logically: the tail is scalar)
__global__ void __launch_bounds__(256) adam_ema_kernel(uint64_t n, uint64_t n_matrix, AdamHyper h, float* __restrict__ w32, __half* __restrict__ w16,
                                                        const __half* __restrict__ g16, float* __restrict__ m1, float* __restrict__ m2,
                                                        uint32_t* __restrict__ steps, __half* __restrict__ ema16) {
	const uint64_t n4 = n / 4;
	for (uint64_t q = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; q < n4; q += (uint64_t)gridDim.x * blockDim.x) {
		const uint64_t i = q * 4;
		const uint2 gp = reinterpret_cast<const uint2*>(g16)[q];
		uint2 wp = reinterpret_cast<const uint2*>(w16)[q];
		const uint2 ep = reinterpret_cast<const uint2*>(ema16)[q];
		float gh[4]; __half wh[4]; float eh[4];
		{
			const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&gp.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&gp.y));
			gh[0] = a.x; gh[1] = a.y; gh[2] = b.x; gh[3] = b.y;
			const __half2 w0 = *reinterpret_cast<const __half2*>(&wp.x), w1 = *reinterpret_cast<const __half2*>(&wp.y);
			wh[0] = __low2half(w0); wh[1] = __high2half(w0); wh[2] = __low2half(w1); wh[3] = __high2half(w1);
			const float2 c = __half22float2(*reinterpret_cast<const __half2*>(&ep.x)), d = __half22float2(*reinterpret_cast<const __half2*>(&ep.y));
			eh[0] = c.x; eh[1] = c.y; eh[2] = d.x; eh[3] = d.y;
		}
		const bool any_matrix = i < n_matrix;
		const bool any = any_matrix || gh[0] != 0.f || gh[1] != 0.f || gh[2] != 0.f || gh[3] != 0.f;
		if (any) {
			float4 w = reinterpret_cast<float4*>(w32)[q], a = reinterpret_cast<float4*>(m1)[q], b = reinterpret_cast<float4*>(m2)[q];
			uint4 s = reinterpret_cast<uint4*>(steps)[q];
			float* wv = &w.x; float* av = &a.x; float* bv = &b.x; uint32_t* sv = &s.x;
			#pragma unroll
			for (int k = 0; k < 4; ++k) { bool t; adam_one(h, i + k < n_matrix, gh[k], wv[k], av[k], bv[k], sv[k], wh[k], t); }
			reinterpret_cast<float4*>(w32)[q] = w; reinterpret_cast<float4*>(m1)[q] = a; reinterpret_cast<float4*>(m2)[q] = b;
			reinterpret_cast<uint4*>(steps)[q] = s;
			const __half2 w0 = __halves2half2(wh[0], wh[1]), w1 = __halves2half2(wh[2], wh[3]);
			wp.x = *reinterpret_cast<const uint32_t*>(&w0); wp.y = *reinterpret_cast<const uint32_t*>(&w1);
			reinterpret_cast<uint2*>(w16)[q] = wp;
		}
		float e[4];
		#pragma unroll
		for (int k = 0; k < 4; ++k) e[k] = (eh[k] * h.ema_decay * h.deb_old + __half2float(wh[k]) * (1 - h.ema_decay)) * h.deb_new;
		const __half2 e0 = __floats2half2_rn(e[0], e[1]), e1 = __floats2half2_rn(e[2], e[3]);
		uint2 eo; eo.x = *reinterpret_cast<const uint32_t*>(&e0); eo.y = *reinterpret_cast<const uint32_t*>(&e1);
		reinterpret_cast<uint2*>(ema16)[q] = eo;
	}
	// scalar tail
	if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
		const uint64_t i = n4 * 4 + threadIdx.x;
		__half wh = w16[i]; bool t;
		adam_one(h, i < n_matrix, __half2float(g16[i]), w32[i], m1[i], m2[i], steps[i], wh, t);
		w16[i] = wh;
		ema16[i] = __float2half_rn((__half2float(ema16[i]) * h.ema_decay * h.deb_old + __half2float(wh) * (1 - h.ema_decay)) * h.deb_new);
	}
}

} // namespace ngp

using namespace ngp;

extern "C" int ngp_init_params(void* stream, uint64_t* rng_state, uint64_t rng_inc, uint64_t n_grid_params, float* d_params_fp32, void* d_params_fp16) {
	NGP_REQUIRE(rng_state && d_params_fp32 && d_params_fp16, "null pointer");
	cudaStream_t s = (cudaStream_t)stream;
	Pcg32 rnd; rnd.state = *rng_state; rnd.inc = rng_inc;
	// density MLP then rgb MLP, xavier uniform drawn one float at a time on the host (gpu_matrix.h:147-164)
	static const uint32_t shapes[5][2] = {{64, 32}, {16, 64}, {64, 32}, {64, 64}, {16, 64}};
	std::vector<float> w(NGP_MLP_PARAMS);
	std::vector<__half> wh(NGP_MLP_PARAMS);
	size_t pos = 0;
	for (int l = 0; l < 5; ++l) {
		const float scale = std::sqrt(6.0f / (float)(shapes[l][0] + shapes[l][1]));
		for (uint32_t i = 0; i < shapes[l][0] * shapes[l][1]; ++i, ++pos) { w[pos] = rnd.next_float() * 2.0f * scale - scale; wh[pos] = __float2half_rn(w[pos]); }
	}
	if (cudaMemcpyAsync(d_params_fp32, w.data(), NGP_MLP_PARAMS * sizeof(float), cudaMemcpyHostToDevice, s) != cudaSuccess ||
	    cudaMemcpyAsync(d_params_fp16, wh.data(), NGP_MLP_PARAMS * sizeof(__half), cudaMemcpyHostToDevice, s) != cudaSuccess ||
	    cudaStreamSynchronize(s) != cudaSuccess) { set_error("ngp_init_params: copy failed: %s", cudaGetErrorString(cudaGetLastError())); return -2; }
	if (n_grid_params) {
		const uint64_t n_threads = (n_grid_params + 3) / 4, total_threads = ((n_threads + 127) / 128) * 128;
		NGP_LAUNCH("init_grid_kernel", s, init_grid_kernel<<<grid_for(n_threads, 256), 256, 0, s>>>(n_grid_params, n_threads, total_threads, rnd, d_params_fp32 + NGP_MLP_PARAMS, (__half*)d_params_fp16 + NGP_MLP_PARAMS));
	}
	rnd.advance(n_grid_params);
	*rng_state = rnd.state;
	return check_launch("init_params");
}

extern "C" int ngp_optimizer_step(void* stream, const ngp_optimizer_cfg* cfg, uint64_t n_params, uint64_t n_matrix_params, uint32_t step, float learning_rate,
                                  float* d_w32, void* d_w16, const void* d_g16, float* d_m1, float* d_m2, uint32_t* d_param_steps, void* d_ema16) {
	NGP_REQUIRE(cfg && d_w32 && d_w16 && d_g16 && d_m1 && d_m2 && d_param_steps && d_ema16, "null pointer");
	NGP_REQUIRE(n_matrix_params % 4 == 0, "n_matrix_params must be a multiple of 4");
	if (n_params == 0) return 0;
	const uint32_t new_step = step + 1;
	AdamHyper h;
	h.lr = learning_rate; h.beta1 = cfg->beta1; h.beta2 = cfg->beta2; h.eps = cfg->epsilon; h.l2_reg = cfg->l2_reg; h.loss_scale = cfg->loss_scale;
	h.ema_decay = cfg->ema_decay;
	h.deb_old = 1 - (float)std::pow(cfg->ema_decay, new_step - 1);
	h.deb_new = 1.0f / (1 - (float)std::pow(cfg->ema_decay, new_step));
	NGP_LAUNCH("adam_ema_kernel", (cudaStream_t)stream, adam_ema_kernel<<<grid_for(n_params / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>(n_params, n_matrix_params, h, d_w32, (__half*)d_w16, (const __half*)d_g16,
		d_m1, d_m2, d_param_steps, (__half*)d_ema16));
	return check_launch("optimizer_step");
}
