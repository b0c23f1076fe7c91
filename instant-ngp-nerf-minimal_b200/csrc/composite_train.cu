// Volume-render compositing for training: forward, loss, analytic backward, ordered compaction.  sm_100a.
// Replaces volume_rendering_with_loss_and_gradient_kernal (internal/render/ray_marcher.cu:15-213) and the two rollover
// kernels (tiny-cuda-nn/common_device.h:485-505).  The reference walks each ray with one thread, three times; here a WARP
// owns a ray and the running transmittance / prefix colour are per-ray segmented warp scans, so a step's ~8 k rays become
// ~260 k busy lanes.  The compacted base of each ray is an exclusive scan in ray order (bit-reproducible) instead of one
// global atomicAdd per ray.  HBM-bound: 12 B / marched sample in the forward launch, 72 B / trained sample in the backward one.
#include "composite.cuh"
#include <mutex>

namespace ngp {

// per-ray scratch between the three launches
struct RayTmp { float r, g, b, T; };

__device__ __forceinline__ float warp_scan_mul(float v, uint32_t lane) {
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) { const float t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v *= t; }
	return v;
}
__device__ __forceinline__ float warp_scan_add(float v, uint32_t lane) {
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) { const float t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
	return v;
}
__device__ __forceinline__ float warp_sum(float v) {
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}

// Launch 1: forward compositing, ONE WARP PER RAY.  Lanes take 32 consecutive samples; the transmittance in front of each
// sample is a segmented (per-ray) multiplicative warp scan carried from chunk to chunk; the first sample whose transmittance
// fell below the threshold ends the ray (ballot + ffs), exactly where the reference's sequential loop breaks.
__global__ void __launch_bounds__(256) composite_count_kernel(
	ngp_render_cfg c, uint32_t n_rays, const uint32_t* __restrict__ ray_counter, const uint32_t* __restrict__ numsteps_in,
	const Coordinate* __restrict__ coords_in, const __half* __restrict__ net_out, RayTmp* __restrict__ tmp, uint32_t* __restrict__ n_used_out)
{
	chain_enter();
	const uint32_t lane = threadIdx.x & 31, i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const uint32_t n_kept = min(*ray_counter, n_rays);
	if (i >= n_kept) return;
	const uint32_t numst = numsteps_in[i * 2], base = numsteps_in[i * 2 + 1];
	const Coordinate* cs = coords_in + base;
	const __half* no = net_out + (size_t)base * 4;
	float T = 1.f, acc[3] = {0.f, 0.f, 0.f};
	uint32_t n_used = numst;
	// The kernel lasts as long as its longest ray (every warp is resident at once), i.e. chunks x the latency of one chunk: the
	// loads of chunk c + 1 are requested before chunk c is scanned, so a chunk costs its shuffles, not a trip to L2 / HBM.
	uint2 nx_raw = make_uint2(0u, 0u); float nx_dt = 0.f;
	if (lane < numst) { nx_raw = *reinterpret_cast<const uint2*>(no + (size_t)lane * 4); nx_dt = cs[lane].dt; }
	for (uint32_t c0 = 0; c0 < numst; c0 += 32) {
		const uint32_t k = c0 + lane;
		const bool valid = k < numst;
		const uint2 cur_raw = nx_raw; const float cur_dt = nx_dt;
		if (k + 32 < numst) { nx_raw = *reinterpret_cast<const uint2*>(no + (size_t)(k + 32) * 4); nx_dt = cs[k + 32].dt; }
		float raw[4] = {0.f, 0.f, 0.f, 0.f}, alpha = 0.f;
		if (valid) {
			unpack_rgbsigma(cur_raw, raw);
			const float dt = unwarp_dt(cur_dt, c.min_step, c.n_cascades);
			alpha = 1.f - __expf(-net_to_density(raw[3], c.density_activation) * dt);
		}
		const float incl = warp_scan_mul(1.f - alpha, lane);
		float excl = __shfl_up_sync(0xffffffffu, incl, 1);
		if (lane == 0) excl = 1.f;
		const float Tk = T * excl;                               // transmittance in front of sample k
		const uint32_t stop = __ballot_sync(0xffffffffu, valid && Tk < c.transmittance_threshold);
		const uint32_t first = stop ? (uint32_t)__ffs(stop) - 1u : 32u;
		if (valid && lane < first) {
			const float w = alpha * Tk;
			acc[0] += w * net_to_rgb(raw[0], c.rgb_activation);
			acc[1] += w * net_to_rgb(raw[1], c.rgb_activation);
			acc[2] += w * net_to_rgb(raw[2], c.rgb_activation);
		}
		if (stop) { n_used = c0 + first; break; }
		T *= __shfl_sync(0xffffffffu, incl, 31);
	}
	const float r = warp_sum(acc[0]), g = warp_sum(acc[1]), b = warp_sum(acc[2]);
	if (lane == 0) { tmp[i] = {r, g, b, T}; n_used_out[i] = n_used; }
}

// Launch 2: exclusive prefix sum of the per-ray sample counts in ray order (one CTA: at most 2^18 rays, a few microseconds).
// Replaces one global atomicAdd per ray and makes the compacted batch bit-reproducible.
__global__ void __launch_bounds__(1024) composite_scan_kernel(uint32_t n_rays, const uint32_t* __restrict__ ray_counter, const uint32_t* __restrict__ n_used,
                                                              uint32_t* __restrict__ cbase, uint32_t* __restrict__ compacted_counter) {
	chain_enter();
	__shared__ uint32_t warp_tot[32];
	const uint32_t n_kept = min(*ray_counter, n_rays);
	const uint32_t per = (n_kept + 1023) / 1024, lo = min(threadIdx.x * per, n_kept), hi = min(lo + per, n_kept);
	uint32_t sum = 0;
	for (uint32_t j = lo; j < hi; ++j) sum += n_used[j];
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t incl = sum;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
	if (lane == 31) warp_tot[warp] = incl;
	__syncthreads();
	if (warp == 0) {
		const uint32_t w = warp_tot[lane];
		uint32_t wi = w;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += t; }
		warp_tot[lane] = wi - w;
		if (lane == 31) *compacted_counter = wi;      // final value, may exceed the batch like the reference's atomicAdd total
	}
	__syncthreads();
	uint32_t run = warp_tot[warp] + incl - sum;
	for (uint32_t j = lo; j < hi; ++j) { cbase[j] = run; run += n_used[j]; }
}

// Launch 3: loss + analytic backward + compaction, one warp per ray.  Prefix colour and transmittance come from warp scans;
// per-sample gradients are written as coalesced 8-byte records, the ray's Coordinates are copied as one flat float run.
__global__ void __launch_bounds__(256) composite_grad_kernel(
	ngp_render_cfg c, uint32_t n_rays, uint32_t max_compacted, const uint32_t* __restrict__ ray_counter, const float4* __restrict__ rays_rgba,
	uint32_t* __restrict__ numsteps_io, const Coordinate* __restrict__ coords_in, Coordinate* __restrict__ coords_compacted,
	const __half* __restrict__ net_out, __half* __restrict__ dloss, uint32_t padded_width, const float* __restrict__ mean_density_ptr, Pcg32 rng,
	float* __restrict__ loss_out, const RayTmp* __restrict__ tmp, const uint32_t* __restrict__ n_used_in, const uint32_t* __restrict__ cbase_in,
	uint32_t* __restrict__ src_index)
{
	chain_enter();
	const uint32_t lane = threadIdx.x & 31, i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	const uint32_t n_kept = min(*ray_counter, n_rays);
	if (i >= n_kept) return;
	const uint32_t numst = numsteps_io[i * 2], base = numsteps_io[i * 2 + 1];
	const uint32_t n_used = n_used_in[i], cbase = cbase_in[i];
	const RayTmp t0 = tmp[i];
	float rgb_ray[3] = {t0.r, t0.g, t0.b};

	// background: same random colour for every ray of a step (un-advanced generator, ray_marcher.cu:90-93)
	float bg[3] = {c.background_color[0], c.background_color[1], c.background_color[2]};
	if (c.train_with_random_bg_color) { bg[0] = rng.next_float(); bg[1] = rng.next_float(); bg[2] = rng.next_float(); }
	#pragma unroll
	for (int k = 0; k < 3; ++k) bg[k] = srgb_to_linear(bg[k]);
	// the environment map sits between the volume and the flat background (ray_marcher.cu:96-102)
	float dir[3] = {0.f, 0.f, 0.f};
	if (c.envmap) {
		dir[0] = c.rays[(size_t)i * 6 + 3]; dir[1] = c.rays[(size_t)i * 6 + 4]; dir[2] = c.rays[(size_t)i * 6 + 5];
		float ev[4];
		read_envmap(reinterpret_cast<const __half*>(c.envmap), c.envmap_res_x, c.envmap_res_y, dir[0], dir[1], dir[2], ev);
		#pragma unroll
		for (int k = 0; k < 3; ++k) bg[k] = ev[k] + bg[k] * (1.0f - ev[3]);
	}
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
		for (int k = 0; k < 3; ++k) rgb_ray[k] += t0.T * bg[k];
	}

	const uint32_t n = min(max_compacted - min(max_compacted, cbase), n_used);
	__syncwarp();
	if (lane == 0) { numsteps_io[i * 2] = n; numsteps_io[i * 2 + 1] = cbase; }
	if (n == 0) return;

	float loss[3], g[3];
	loss_and_grad(c.loss_type, tgt, rgb_ray, loss, g);
	if (loss_out && lane == 0) loss_out[i] = ((loss[0] + loss[1] + loss[2]) / 3.f) / (float)n_rays;
	const float ls = c.loss_scale / (float)n_rays;
	const float l2reg = c.rgb_activation == ACT_EXP ? 1e-4f : 0.f;
	const float l1reg = *mean_density_ptr < c.min_optical_thickness ? 1e-4f * (c.reg_scale > 0.f ? c.reg_scale : 1.f) : 0.f;

	const Coordinate* cs = coords_in + base;
	const __half* no = net_out + (size_t)base * 4;
	float T = 1.f, pre[3] = {0.f, 0.f, 0.f};          // carried: transmittance and colour in front of the chunk
	uint2 nx_raw = make_uint2(0u, 0u); float nx_dt = 0.f;        // next chunk's loads, requested one chunk ahead (see composite_count_kernel)
	if (lane < n) { nx_raw = *reinterpret_cast<const uint2*>(no + (size_t)lane * 4); nx_dt = cs[lane].dt; }
	for (uint32_t c0 = 0; c0 < n; c0 += 32) {
		const uint32_t k = c0 + lane;
		const bool valid = k < n;
		const uint2 cur_raw = nx_raw; const float cur_dt = nx_dt;
		if (k + 32 < n) { nx_raw = *reinterpret_cast<const uint2*>(no + (size_t)(k + 32) * 4); nx_dt = cs[k + 32].dt; }
		float raw[4] = {0.f, 0.f, 0.f, 0.f}, alpha = 0.f, dt = 0.f;
		if (valid) {
			unpack_rgbsigma(cur_raw, raw);
			dt = unwarp_dt(cur_dt, c.min_step, c.n_cascades);
			alpha = 1.f - __expf(-net_to_density(raw[3], c.density_activation) * dt);
		}
		const float incl = warp_scan_mul(1.f - alpha, lane);
		float excl = __shfl_up_sync(0xffffffffu, incl, 1);
		if (lane == 0) excl = 1.f;
		const float Tk = T * excl, w = alpha * Tk, T_after = Tk * (1.f - alpha);
		const float rgb[3] = {net_to_rgb(raw[0], c.rgb_activation), net_to_rgb(raw[1], c.rgb_activation), net_to_rgb(raw[2], c.rgb_activation)};
		float rgb2[3];
		#pragma unroll
		for (int j = 0; j < 3; ++j) rgb2[j] = pre[j] + warp_scan_add(valid ? w * rgb[j] : 0.f, lane);     // colour up to and including sample k
		if (valid) {
			float dl[4];
			#pragma unroll
			for (int j = 0; j < 3; ++j) dl[j] = ls * (w * g[j] * net_to_rgb_d(raw[j], c.rgb_activation) + fmaxf(0.f, l2reg * raw[j]));
			float dot = 0.f;
			#pragma unroll
			for (int j = 0; j < 3; ++j) dot += g[j] * (T_after * rgb[j] - (rgb_ray[j] - rgb2[j]));
			const float dmlp = net_to_density_d(raw[3], c.density_activation) * (dt * dot);
			dl[3] = ls * dmlp + (raw[3] < 0 ? -l1reg : 0.f);
			const __half2 lo = __floats2half2_rn(dl[0], dl[1]), hi = __floats2half2_rn(dl[2], dl[3]);
			uint2 pk; pk.x = *reinterpret_cast<const uint32_t*>(&lo); pk.y = *reinterpret_cast<const uint32_t*>(&hi);
			*reinterpret_cast<uint2*>(dloss + (size_t)(cbase + k) * padded_width) = pk;
		}
		T *= __shfl_sync(0xffffffffu, incl, 31);
		#pragma unroll
		for (int j = 0; j < 3; ++j) pre[j] = __shfl_sync(0xffffffffu, rgb2[j], 31);
	}
	// gradient of the loss w.r.t. the environment map, only for rays that were traversed to their end (ray_marcher.cu:190-212):
	// the background entered the ray colour with the final transmittance; alpha receives no gradient (:208-209)
	if (c.envmap_gradient && n == numst && lane == 0) {
		float ge[3] = {g[0], g[1], g[2]};
		if (c.envmap_loss_type != c.loss_type) { float l2[3]; loss_and_grad(c.envmap_loss_type, tgt, rgb_ray, l2, ge); }
		float v[4];
		#pragma unroll
		for (int k = 0; k < 3; ++k) {
			float d = t0.T * ge[k];
			if (!c.train_in_linear_color) d /= srgb_to_linear_derivative(bg[k]);
			v[k] = ls * d;
		}
		v[3] = 0.f;
		deposit_envmap_gradient(v, reinterpret_cast<__half*>(c.envmap_gradient), c.envmap_res_x, c.envmap_res_y, dir[0], dir[1], dir[2]);
	}
	// compaction: the ray's n Coordinates are one contiguous run of 7n floats on both sides
	const float* src = reinterpret_cast<const float*>(cs);
	float* dst = reinterpret_cast<float*>(coords_compacted + cbase);
	for (uint32_t j = lane; j < n * 7; j += 32) dst[j] = src[j];
	// where each compacted sample came from, so the training pass can reuse the encodings of the inference pass
	if (src_index) for (uint32_t j = lane; j < n; j += 32) src_index[cbase + j] = base + j;
}

// fill_rollover_and_rescale (grads) + fill_rollover (coords) in one launch
__global__ void rollover_kernel(uint32_t batch, const uint32_t* __restrict__ n_in_ptr, __half* __restrict__ grads, uint32_t stride, Coordinate* __restrict__ coords,
                                uint32_t* __restrict__ src_index) {
	chain_enter();
	const uint32_t n_in = *n_in_ptr;
	if (n_in == 0 || n_in >= batch) return;
	for (uint32_t i = n_in + blockIdx.x * blockDim.x + threadIdx.x; i < batch; i += gridDim.x * blockDim.x) {
		const uint32_t src = i % n_in;
		coords[i] = coords[src];
		if (src_index) src_index[i] = src_index[src];
		for (uint32_t k = 0; k < stride; ++k)
			grads[(size_t)i * stride + k] = __float2half_rn(__half2float(grads[(size_t)src * stride + k]) * (float)n_in / (float)batch);
	}
}

} // namespace ngp

using namespace ngp;

// workspace: [RayTmp n_rays][n_used n_rays][cbase n_rays]
extern "C" size_t ngp_composite_workspace_bytes(uint32_t n_rays) { return (size_t)n_rays * (sizeof(RayTmp) + 8) + 256; }

extern "C" int ngp_composite_train(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays, uint32_t max_compacted, const uint32_t* d_ray_counter,
                                   const float* d_rays_rgba, uint32_t* d_numsteps, const float* d_coords, float* d_coords_compacted,
                                   uint32_t* d_compacted_counter, const void* d_rgbsigma, void* d_dloss, uint32_t padded_width,
                                   const float* d_mean_density, uint64_t rng_state, uint64_t rng_inc, float* d_loss, void* d_scan_ws, uint32_t* d_src_index) {
	NGP_REQUIRE(cfg && d_ray_counter && d_rays_rgba && d_numsteps && d_coords && d_coords_compacted && d_compacted_counter && d_rgbsigma && d_dloss && d_mean_density && d_scan_ws, "null pointer");
	NGP_REQUIRE(padded_width >= 4 && padded_width % 4 == 0, "padded_width must be a multiple of 4");
	NGP_REQUIRE(n_rays > 0, "n_rays must be positive");
	NGP_REQUIRE(!(cfg->envmap || cfg->envmap_gradient) || (cfg->rays && cfg->envmap && cfg->envmap_res_x > 0 && cfg->envmap_res_y > 0), "an environment map needs its data, its resolution and the ray directions");
	cudaStream_t s = (cudaStream_t)stream;
	Pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
	RayTmp* tmp = (RayTmp*)d_scan_ws;
	uint32_t* n_used = (uint32_t*)(tmp + n_rays);
	uint32_t* cbase = n_used + n_rays;
	const uint32_t grid = div_up(n_rays, 8);
	NGP_LAUNCH("composite_count_kernel", s, launch_chained(composite_count_kernel, grid, 256, 0, s, *cfg, n_rays, d_ray_counter, d_numsteps, (const Coordinate*)d_coords,
		(const __half*)d_rgbsigma, tmp, n_used));
	NGP_LAUNCH("composite_scan_kernel", s, launch_chained(composite_scan_kernel, 1, 1024, 0, s, n_rays, d_ray_counter, n_used, cbase, d_compacted_counter));
	NGP_LAUNCH("composite_grad_kernel", s, launch_chained(composite_grad_kernel, grid, 256, 0, s, *cfg, n_rays, max_compacted, d_ray_counter, (const float4*)d_rays_rgba, d_numsteps,
		(const Coordinate*)d_coords, (Coordinate*)d_coords_compacted, (const __half*)d_rgbsigma, (__half*)d_dloss, padded_width, d_mean_density, rng, d_loss,
		tmp, n_used, cbase, d_src_index));
	return check_launch("composite_train");
}

// ---- no-allocation variant (no counterpart in the reference) ----
// Same launches as ngp_composite_train, with the per-ray scratch in a file-scope device array instead of a caller-provided buffer,
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops).  Every GPU that loads the library holds the
// array (1.5 MB) until exit.  In host code the array's name is a host-side stand-in, not a device address: the address is looked
// up once per device with cudaGetSymbolAddress and cached.  All calls share the array: calls that overlap on one GPU race.
static __device__ __align__(256) unsigned char g_composite_scratch[(size_t)NGP_COMPOSITE_STATIC_MAX_RAYS * (sizeof(RayTmp) + 8) + 256];

static void* composite_scratch_address() {
	static std::mutex lock;
	static void* address[64] = {};
	int dev = -1;
	if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("ngp_composite_train_static: cudaGetDevice failed or device index above 63"); return nullptr; }
	std::lock_guard<std::mutex> guard(lock);
	if (!address[dev]) {
		void* p = nullptr;
		const cudaError_t e = cudaGetSymbolAddress(&p, g_composite_scratch);        // with `dev` current: this device's copy of the array
		if (e != cudaSuccess || !p) { set_error("ngp_composite_train_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e)); return nullptr; }
		address[dev] = p;
	}
	return address[dev];
}

extern "C" int ngp_composite_train_static(void* stream, const ngp_render_cfg* cfg, uint32_t n_rays, uint32_t max_compacted, const uint32_t* d_ray_counter,
                                          const float* d_rays_rgba, uint32_t* d_numsteps, const float* d_coords, float* d_coords_compacted,
                                          uint32_t* d_compacted_counter, const void* d_rgbsigma, void* d_dloss, uint32_t padded_width,
                                          const float* d_mean_density, uint64_t rng_state, uint64_t rng_inc, float* d_loss, uint32_t* d_src_index) {
	if (n_rays > NGP_COMPOSITE_STATIC_MAX_RAYS) {
		set_error("ngp_composite_train_static: n_rays = %u exceeds the static scratch (NGP_COMPOSITE_STATIC_MAX_RAYS = %u); use ngp_composite_train with a caller-owned workspace", n_rays, NGP_COMPOSITE_STATIC_MAX_RAYS);
		return -1;
	}
	static_assert(sizeof(g_composite_scratch) >= (size_t)NGP_COMPOSITE_STATIC_MAX_RAYS * (sizeof(RayTmp) + 8) + 256, "scratch covers the maximum");
	void* ws = composite_scratch_address();
	if (!ws) return -2;
	// nothing in the scratch has to start at zero: composite_count_kernel writes every slot the later launches read
	return ngp_composite_train(stream, cfg, n_rays, max_compacted, d_ray_counter, d_rays_rgba, d_numsteps, d_coords, d_coords_compacted, d_compacted_counter, d_rgbsigma,
	                           d_dloss, padded_width, d_mean_density, rng_state, rng_inc, d_loss, ws, d_src_index);
}

extern "C" int ngp_rollover(void* stream, uint32_t batch, const uint32_t* d_compacted_counter, void* d_dloss, uint32_t padded_width, float* d_coords_compacted, uint32_t* d_src_index) {
	NGP_REQUIRE(d_compacted_counter && d_dloss && d_coords_compacted, "null pointer");
	NGP_LAUNCH("rollover_kernel", (cudaStream_t)stream, launch_chained(rollover_kernel, grid_for(batch, 256), 256, 0, (cudaStream_t)stream, batch, d_compacted_counter, (__half*)d_dloss, padded_width, (Coordinate*)d_coords_compacted, d_src_index));
	return check_launch("rollover");
}
