// Parameter initialisation and the fused Adam + EMA step for sm_100a.
// Replaces Trainer::initialize_params (tiny-cuda-nn/trainer.h:73-112), generate_random_kernel (random.h:69-94),
// adam_step (optimizers/adam.h:50-113) and ema_step_half_precision (optimizers/ema.h:65-79): one pass over the arena
// instead of two, 16-byte vector accesses, and the EMA of an untouched grid entry is still updated (it decays towards the
// unchanged weight exactly like the reference).  HBM-bound: 42 B / parameter when touched, 6 B when skipped.
#include "common.cuh"
#include <vector>
#include <stdlib.h>
#include <cmath>

namespace ngp {

// reference pattern: thread i advances the generator by 4*i and writes elements i + total_threads*j, j < 4; ALL launched
// threads take part (total_threads = div_round_up(n, 4) rounded up to whole 128-thread blocks), verified against a dump
__global__ void init_grid_kernel(uint64_t n, uint64_t total_threads, Pcg32 rng0, float* __restrict__ w32, __half* __restrict__ w16) {
	for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total_threads; i += (uint64_t)gridDim.x * blockDim.x) {
		Pcg32 rng = rng0;
		rng.advance(i * 4);
		for (uint64_t j = 0; j < 4; ++j) {
			const uint64_t idx = i + total_threads * j;
			if (idx >= n) break;
			const float v = __fmaf_rn(rng.next_float(), 1e-4f - (-1e-4f), -1e-4f);     // the reference's lambda is contracted to one FFMA
			w32[idx] = v;
			w16[idx] = __float2half_rn(v);
		}
	}
}

struct AdamHyper {
	float lr, beta1, beta2, eps, l2_reg, loss_scale, ema_decay, deb_old, deb_new;
	float inv_loss_scale;                   // exact reciprocal when loss_scale is a power of two, else 0 (then the kernel divides)
	float l2b1_hi, l2b1_lo, l2b2_hi, l2b2_lo; // log2(beta) as an unevaluated fp32 sum, for beta^steps = exp2(steps * log2 beta)
};

// beta^c for an integer-valued c: the product c*log2(beta) is formed to ~48 bits (exact head product via fma, plus the
// tail), so the result is as accurate as powf (a few ulp) at a tenth of its instruction count.
__device__ __forceinline__ float pow_steps(float c, float hi, float lo) {
	const float p = c * hi;
	const float e = __fmaf_rn(c, hi, -p) + c * lo;
	const float r = exp2f(p);
	return __fmaf_rn(r, e * 0.693147180559945f, r);
}

// lr * sqrt(1 - beta2^cs) / (1 - beta1^cs): the per-parameter bias correction of adam.h:85-90
__device__ __forceinline__ float bias_corrected_lr(const AdamHyper& h, uint32_t cs) {
	const float c = (float)cs;
	return h.lr * (sqrtf(1 - pow_steps(c, h.l2b2_hi, h.l2b2_lo)) / (1 - pow_steps(c, h.l2b1_hi, h.l2b1_lo)));
}

// cached_steps / cached_lr: neighbouring parameters usually share their step count, so the correction is computed once per run
__device__ __forceinline__ void adam_one(const AdamHyper& h, bool matrix, float gh, float& w, float& m1, float& m2, uint32_t& steps, __half& w16, bool& touched,
                                         uint32_t& cached_steps, float& cached_lr) {
	float g = h.inv_loss_scale != 0.f ? gh * h.inv_loss_scale : gh / h.loss_scale;
	touched = matrix || g != 0.f;
	if (!touched) return;
	if (matrix) g += h.l2_reg * w;
	const float gsq = g * g;
	m1 = h.beta1 * m1 + (1 - h.beta1) * g;
	m2 = h.beta2 * m2 + (1 - h.beta2) * gsq;
	const uint32_t cs = ++steps;
	if (cs != cached_steps) { cached_steps = cs; cached_lr = bias_corrected_lr(h, cs); }
	const float lr = cached_lr;
	const float eff = fminf(fmaxf(lr / (sqrtf(m2) + h.eps), 0.f), 3.402823466e38f);
	w = w - eff * m1;
	w16 = __float2half_rn(w);
}

// system-scope flag accesses and peer loads of the data-parallel step (peers' memory is mapped through CUDA IPC)
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) { uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }
// peer loads must come from the owner's L2, never from a stale local line
__device__ __forceinline__ uint4 ld_peer16(const void* p) { uint4 v; asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory"); return v; }

__device__ __forceinline__ void unpack8(const uint4& p, float* f) {
	const uint32_t w[4] = {p.x, p.y, p.z, p.w};
	#pragma unroll
	for (int i = 0; i < 4; ++i) { const float2 t = __half22float2(*reinterpret_cast<const __half2*>(&w[i])); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
	uint32_t w[4];
	#pragma unroll
	for (int i = 0; i < 4; ++i) { const __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]); w[i] = *reinterpret_cast<const uint32_t*>(&h); }
	return make_uint4(w[0], w[1], w[2], w[3]);
}

// 8 parameters per thread: three 16-byte loads (gradient, fp16 weight, EMA) are issued together; the 64 bytes of fp32
// state are only touched when one of the 8 gradients is non-zero (or the group belongs to the MLP matrices).
// PEER: the range [begin, n) is this rank's shard; new fp16 weights are stored into every rank's arena (all-gather half of the
// sharded step) and the last CTA to finish raises the "weights of this epoch have landed" flag at every rank.
// PEER with WORLD > 0: the reduce-scatter is fused in -- the kernel first waits for every rank's "gradients complete" flag, then
// each group of 8 gradients is summed straight out of all ranks' arenas (16-byte P2P loads over NVLink, fixed rank order, fp32,
// rounded to fp16 once) while the fp32 state streams from local HBM.
template <bool PEER, int WORLD = 0>
__global__ void __launch_bounds__(256, 3) adam_ema_kernel(uint64_t begin, uint64_t n, uint64_t n_matrix, AdamHyper h, float* __restrict__ w32, __half* __restrict__ w16,
                                                        const __half* g16, float* __restrict__ m1, float* __restrict__ m2,
                                                        uint32_t* __restrict__ steps, __half* __restrict__ ema16, ngp_peer_set ps, uint32_t epoch,
                                                        uint32_t* __restrict__ ticket) {
	chain_enter();
	if (PEER && WORLD > 0) {
		if (threadIdx.x < ps.world) { while (ld_acquire_sys(ps.flags[ps.rank] + threadIdx.x) < epoch) __nanosleep(32); }
		__syncthreads();
	}
	const uint64_t n8 = n / 8;
	for (uint64_t q = begin / 8 + blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; q < n8; q += (uint64_t)gridDim.x * blockDim.x) {
		const uint64_t i = q * 8;
		// (streaming .cs loads/stores for the once-per-step state were measured: the kernel slows from 86 to 104 us and the next
		// step's gathers gain 1 us -- plain policy kept)
		uint4 gp;
		if (PEER && WORLD > 0) {
			// (requesting the NEXT group's peer gradients one iteration ahead was measured: 104-121 registers, one CTA fewer per SM,
			// 88 -> 111 us; the three resident CTAs per SM already cover the NVLink round trip)
			// at most four 16-byte peer loads in flight per thread (8 ranks: two rounds), summed in rank order
			constexpr int kRound = WORLD > 4 ? 4 : (WORLD > 0 ? WORLD : 1);
			float sum[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
			#pragma unroll
			for (int p0 = 0; p0 < WORLD; p0 += kRound) {
				uint4 pg[kRound];
				#pragma unroll
				for (int p = 0; p < kRound; ++p) if (p0 + p < (int)ps.world) pg[p] = ld_peer16(reinterpret_cast<const uint4*>(ps.grads[p0 + p]) + q);
				#pragma unroll
				for (int p = 0; p < kRound; ++p) {
					if (p0 + p < (int)ps.world) {
						float t[8];
						unpack8(pg[p], t);
						#pragma unroll
						for (int k = 0; k < 8; ++k) sum[k] += t[k];
					}
				}
			}
			gp = pack8(sum);
		} else {
			gp = __ldg(reinterpret_cast<const uint4*>(g16) + q);
		}
		const uint4 wp = reinterpret_cast<const uint4*>(w16)[q];
		const uint4 ep = reinterpret_cast<const uint4*>(ema16)[q];
		const bool any = i < n_matrix || ((gp.x | gp.y | gp.z | gp.w) & 0x7fff7fffu) != 0u;      // any gradient that is not +-0
		float wh[8];
		unpack8(wp, wh);
		if (any) {
			float gh[8];
			unpack8(gp, gh);
			float4 w[2], a[2], b[2]; uint4 st[2];
			#pragma unroll
			for (int v = 0; v < 2; ++v) {
				w[v] = reinterpret_cast<const float4*>(w32)[q * 2 + v]; a[v] = reinterpret_cast<const float4*>(m1)[q * 2 + v];
				b[v] = reinterpret_cast<const float4*>(m2)[q * 2 + v]; st[v] = reinterpret_cast<const uint4*>(steps)[q * 2 + v];
			}
			float* wv = &w[0].x; float* av = &a[0].x; float* bv = &b[0].x; uint32_t* sv = &st[0].x;
			uint32_t cached_steps = 0; float cached_lr = 0.f;
			#pragma unroll
			for (int k = 0; k < 8; ++k) {
				bool t; __half nh = __float2half_rn(wh[k]);
				adam_one(h, i + k < n_matrix, gh[k], wv[k], av[k], bv[k], sv[k], nh, t, cached_steps, cached_lr);
				if (t) wh[k] = __half2float(nh);
			}
			#pragma unroll
			for (int v = 0; v < 2; ++v) {
				reinterpret_cast<float4*>(w32)[q * 2 + v] = w[v]; reinterpret_cast<float4*>(m1)[q * 2 + v] = a[v];
				reinterpret_cast<float4*>(m2)[q * 2 + v] = b[v]; reinterpret_cast<uint4*>(steps)[q * 2 + v] = st[v];
			}
			const uint4 nw = pack8(wh);
			if (PEER) {
				#pragma unroll 1
				for (uint32_t p = 0; p < ps.world; ++p) reinterpret_cast<uint4*>(ps.params[p])[q] = nw;
			} else {
				reinterpret_cast<uint4*>(w16)[q] = nw;
			}
		}
		float eh[8], e[8];
		unpack8(ep, eh);
		#pragma unroll
		for (int k = 0; k < 8; ++k) e[k] = (eh[k] * h.ema_decay * h.deb_old + wh[k] * (1 - h.ema_decay)) * h.deb_new;
		reinterpret_cast<uint4*>(ema16)[q] = pack8(e);
	}
	if (PEER) {
		__threadfence_system();
		__syncthreads();
		__shared__ bool last;
		if (threadIdx.x == 0) last = atomicAdd(ticket, 1u) == gridDim.x - 1;
		__syncthreads();
		if (last) {
			if (threadIdx.x == 0) *ticket = 0;
			__threadfence_system();
			if (threadIdx.x < ps.world) st_release_sys(ps.flags[threadIdx.x] + NGP_MAX_PEERS + ps.rank, epoch);
		}
		return;
	}
	// scalar tail
	if (blockIdx.x == 0 && threadIdx.x < (n & 7)) {
		const uint64_t i = n8 * 8 + threadIdx.x;
		__half wh = w16[i]; bool t; uint32_t cs = 0; float clr = 0.f;
		adam_one(h, i < n_matrix, __half2float(g16[i]), w32[i], m1[i], m2[i], steps[i], wh, t, cs, clr);
		w16[i] = wh;
		ema16[i] = __float2half_rn((__half2float(ema16[i]) * h.ema_decay * h.deb_old + __half2float(wh) * (1 - h.ema_decay)) * h.deb_new);
	}
}

// Free-running trace of the data-parallel step: the publish kernel stamps "my backward pass is complete", the wait kernel
// stamps "all shards have landed" (globaltimer, ns, slot = epoch mod 4096).  Two stores per step; read with ngp_peer_stamps_read.
__device__ unsigned long long g_peer_stamps[2][4096];
__device__ __forceinline__ unsigned long long global_ns() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }

// ---- sharded step over peer memory ---------------------------------------------------------------------------------
__global__ void peer_publish_kernel(ngp_peer_set ps, uint32_t slot_base, uint32_t epoch) {
	chain_enter();
	// stream order has completed this rank's backward pass before this grid started; the release store (a system-scope fence
	// followed by the store) publishes it cumulatively and raises the flag at every rank
	if (threadIdx.x == 0) g_peer_stamps[0][epoch & 4095u] = global_ns();
	if (threadIdx.x < ps.world) st_release_sys(ps.flags[threadIdx.x] + slot_base + ps.rank, epoch);
}

__global__ void peer_wait_kernel(ngp_peer_set ps, uint32_t slot_base, uint32_t epoch) {
	chain_enter();
	if (threadIdx.x < ps.world) { while (ld_acquire_sys(ps.flags[ps.rank] + slot_base + threadIdx.x) < epoch) __nanosleep(64); }
	__syncwarp();
	if (threadIdx.x == 0) g_peer_stamps[1][epoch & 4095u] = global_ns();
}

} // namespace ngp

using namespace ngp;

static void fill_hyper(AdamHyper& h, const ngp_optimizer_cfg* cfg, uint32_t step, float learning_rate);

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
		NGP_LAUNCH("init_grid_kernel", s, init_grid_kernel<<<grid_for(total_threads, 256), 256, 0, s>>>(n_grid_params, total_threads, rnd, d_params_fp32 + NGP_MLP_PARAMS, (__half*)d_params_fp16 + NGP_MLP_PARAMS));
	}
	rnd.advance(n_grid_params);
	*rng_state = rnd.state;
	return check_launch("init_params");
}

extern "C" int ngp_optimizer_step(void* stream, const ngp_optimizer_cfg* cfg, uint64_t n_params, uint64_t n_matrix_params, uint32_t step, float learning_rate,
                                  float* d_w32, void* d_w16, const void* d_g16, float* d_m1, float* d_m2, uint32_t* d_param_steps, void* d_ema16) {
	NGP_REQUIRE(cfg && d_w32 && d_w16 && d_g16 && d_m1 && d_m2 && d_param_steps && d_ema16, "null pointer");
	NGP_REQUIRE(n_matrix_params % 8 == 0, "n_matrix_params must be a multiple of 8");
	if (n_params == 0) return 0;
	AdamHyper h;
	fill_hyper(h, cfg, step, learning_rate);
	NGP_LAUNCH("adam_ema_kernel", (cudaStream_t)stream, launch_chained(adam_ema_kernel<false>, grid_for(n_params / 8 + 1, 256, 6), 256, 0, (cudaStream_t)stream, 0, n_params, n_matrix_params, h, d_w32, (__half*)d_w16, (const __half*)d_g16,
		d_m1, d_m2, d_param_steps, (__half*)d_ema16, ngp_peer_set{}, 0, nullptr));
	return check_launch("optimizer_step");
}

extern "C" int ngp_peer_stamps_read(uint64_t* host_out /* [2][4096] */) {
	NGP_REQUIRE(host_out, "null pointer");
	if (cudaDeviceSynchronize() != cudaSuccess || cudaMemcpyFromSymbol(host_out, g_peer_stamps, sizeof(unsigned long long) * 2 * 4096) != cudaSuccess) {
		set_error("ngp_peer_stamps_read: %s", cudaGetErrorString(cudaGetLastError())); return -2;
	}
	return 0;
}

static int check_peers(const ngp_peer_set* ps) {
	if (!ps || ps->world < 1 || ps->world > NGP_MAX_PEERS || ps->rank >= ps->world) { set_error("peer set: world must be in [1,%d] and rank < world", NGP_MAX_PEERS); return -1; }
	for (uint32_t p = 0; p < ps->world; ++p) if (!ps->grads[p] || !ps->params[p] || !ps->flags[p]) { set_error("peer set: null pointer for rank %u", p); return -1; }
	return 0;
}

extern "C" int ngp_peer_publish_gradients(void* stream, const ngp_peer_set* peers, uint32_t epoch) {
	if (check_peers(peers)) return -1;
	NGP_LAUNCH("peer_publish_kernel", (cudaStream_t)stream, launch_chained(peer_publish_kernel, 1, 32, 0, (cudaStream_t)stream, *peers, 0, epoch));
	return check_launch("peer_publish_gradients");
}

extern "C" int ngp_peer_wait_weights(void* stream, const ngp_peer_set* peers, uint32_t epoch) {
	if (check_peers(peers)) return -1;
	NGP_LAUNCH("peer_wait_kernel", (cudaStream_t)stream, launch_chained(peer_wait_kernel, 1, 32, 0, (cudaStream_t)stream, *peers, NGP_MAX_PEERS, epoch));
	return check_launch("peer_wait_weights");
}

extern "C" int ngp_optimizer_step_sharded(void* stream, const ngp_optimizer_cfg* cfg, uint64_t shard_begin, uint64_t shard_end, uint64_t n_matrix_params, uint32_t step,
                                          float learning_rate, float* d_w32, float* d_m1, float* d_m2, uint32_t* d_param_steps, void* d_ema16,
                                          const ngp_peer_set* peers, uint32_t epoch, uint32_t* d_ticket) {
	NGP_REQUIRE(cfg && d_w32 && d_m1 && d_m2 && d_param_steps && d_ema16 && d_ticket, "null pointer");
	if (check_peers(peers)) return -1;
	NGP_REQUIRE(shard_begin % 8 == 0 && shard_end % 8 == 0 && shard_begin <= shard_end && n_matrix_params % 8 == 0, "shard bounds and n_matrix_params must be multiples of 8");
	NGP_REQUIRE(epoch >= 1, "epoch starts at 1");
	AdamHyper h;
	fill_hyper(h, cfg, step, learning_rate);
	const uint64_t groups = (shard_end - shard_begin) / 8;
	cudaStream_t st = (cudaStream_t)stream;
	__half* w16 = (__half*)peers->params[peers->rank];
	const __half* g16 = (const __half*)peers->grads[peers->rank];
	const uint32_t agrid = grid_for(groups ? groups : 1, 256, 6);
	// (a separate reduce kernel followed by a plain sharded Adam was measured at 2 GPUs: 0.65 vs 0.625 ms/step fused -- DESIGN.md 7)
	if (peers->world <= 2) {
		NGP_LAUNCH("adam_ema_kernel", st, launch_chained(adam_ema_kernel<true, 2>, agrid, 256, 0, st, shard_begin, shard_end, n_matrix_params, h, d_w32, w16, g16,
			d_m1, d_m2, d_param_steps, (__half*)d_ema16, *peers, epoch, d_ticket));
	} else if (peers->world <= 4) {
		NGP_LAUNCH("adam_ema_kernel", st, launch_chained(adam_ema_kernel<true, 4>, agrid, 256, 0, st, shard_begin, shard_end, n_matrix_params, h, d_w32, w16, g16,
			d_m1, d_m2, d_param_steps, (__half*)d_ema16, *peers, epoch, d_ticket));
	} else {
		NGP_LAUNCH("adam_ema_kernel", st, launch_chained(adam_ema_kernel<true, 8>, agrid, 256, 0, st, shard_begin, shard_end, n_matrix_params, h, d_w32, w16, g16,
			d_m1, d_m2, d_param_steps, (__half*)d_ema16, *peers, epoch, d_ticket));
	}
	return check_launch("optimizer_step_sharded");
}

static void fill_hyper(AdamHyper& h, const ngp_optimizer_cfg* cfg, uint32_t step, float learning_rate) {
	const uint32_t new_step = step + 1;
	h.lr = learning_rate; h.beta1 = cfg->beta1; h.beta2 = cfg->beta2; h.eps = cfg->epsilon; h.l2_reg = cfg->l2_reg; h.loss_scale = cfg->loss_scale;
	h.ema_decay = cfg->ema_decay;
	{
		int ex = 0;
		const bool pow2 = cfg->loss_scale > 0.f && std::isfinite(cfg->loss_scale) && std::frexp(cfg->loss_scale, &ex) == 0.5f && ex > -120 && ex < 120;
		h.inv_loss_scale = pow2 ? 1.0f / cfg->loss_scale : 0.f;
		auto split = [](float beta, float& hi, float& lo) {
			if (!(beta > 0.f)) { hi = -1e20f; lo = 0.f; return; }       // beta^cs == 0 for cs >= 1
			const double l = std::log2((double)beta);
			hi = (float)l; lo = (float)(l - (double)hi);
		};
		split(cfg->beta1, h.l2b1_hi, h.l2b1_lo);
		split(cfg->beta2, h.l2b2_hi, h.l2b2_lo);
	}
	h.deb_old = 1 - (float)std::pow(cfg->ema_decay, new_step - 1);
	h.deb_new = 1.0f / (1 - (float)std::pow(cfg->ema_decay, new_step));
}
