// Training ray generation + occupancy march + ordered compaction, for sm_100a.
// Replaces generate_training_samples_nerf (internal/sampler/occupancy_sampler.cu:310-449).
//
// The reference gives each ray ONE thread that walks the occupancy grid twice (count, then emit) and reserves output space
// with two global atomicAdds; a training step only has ~8 k rays, so the GPU runs a few hundred divergent warps.  Here a WARP
// owns a ray.  The key observation is that the ray parameters the reference visits all lie on one lattice
// t_{k+1} = t_k + calc_dt(t_k) that does not depend on the grid: the occupancy only decides which lattice points are tested.
// So the warp takes the lattice 32 points at a time, every lane evaluates its point in parallel (position, cascade, bitfield
// bit, distance to the next voxel), and a ballot-driven resolve replays the reference's control flow exactly:
//   * an occupied visited point emits a sample and makes the next lattice point visited;
//   * an empty visited point at t with exit distance s makes the next visited point the first later one with t' >= t + s.
// Per-ray counts and coordinates therefore stay bit-identical to the sequential march (same expressions, same order), but the
// bitfield loads and the per-point arithmetic of 32 points overlap, and the resolve costs one ballot per run / per empty voxel.
// Output bases come from a prefix sum in ray order (deterministic), not from atomics.
#include "march.cuh"

namespace ngp {

struct RaySetup { V3 o, d, idir; float startt, u, v; uint32_t img; };

__device__ __forceinline__ RaySetup setup_ray(const ngp_sampler_cfg& c, const ngp_camera_cfg& cam, uint32_t i, uint32_t n_rays, uint32_t n_rays_shift, uint32_t n_images,
                                              const float* __restrict__ xforms, int snap, Pcg32 rng) {
	RaySetup r;
	r.img = (((n_rays_shift + i) * n_images) / n_rays) % n_images;
	rng.advance((uint64_t)i * 8);
	float u = rng.next_float(), v = rng.next_float();
	if (snap) {
		u = ((float)min(max((int)(u * (float)cam.width), 0), cam.width - 1) + 0.5f) / (float)cam.width;
		v = ((float)min(max((int)(v * (float)cam.height), 0), cam.height - 1) + 0.5f) / (float)cam.height;
	}
	r.u = u; r.v = v;
	make_ray(xforms + (size_t)r.img * 12, u, v, cam, r.o, r.d);
	const float tmin = fmaxf(ray_intersect_tmin(c, r.o, r.d), c.near_distance);
	r.startt = __fmaf_rn(calc_dt(tmin, c), rng.next_float(), tmin);
	r.idir = {1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z};
	return r;
}

// One window of the lattice: returns this lane's t (the value after `lane` steps from t_base) and advances t_base by 32 steps.
// With a constant step (cone_angle == 0) the sequentially rounded sums t_{k+1} = fl(t_k + dt) form, inside one binade of t, an
// arithmetic progression of the rounded step dtq = fl(t + dt) - t.  Every lane builds its candidate t_base + lane*dtq and checks
// its own link fl(cand_L + dt) == cand_{L+1}; if all 32 links hold the candidates ARE the sequential values (induction from
// cand_0 == t_base), otherwise (binade crossing, ties, cone stepping) the warp falls back to the sequential recurrence.
__device__ __forceinline__ float lattice_window(float& t_base, uint32_t lane, const ngp_sampler_cfg& c, bool const_dt) {
	if (const_dt) {
		const float dt0 = calc_dt(t_base, c);
		const float dtq = (t_base + dt0) - t_base;
		const float cand = __fmaf_rn((float)lane, dtq, t_base);
		const float next = __fmaf_rn((float)(lane + 1), dtq, t_base);
		if (__all_sync(0xffffffffu, cand + dt0 == next)) { t_base = __shfl_sync(0xffffffffu, next, 31); return cand; }
	}
	float t = t_base, my_t = t_base;
	#pragma unroll
	for (int s = 0; s < 32; ++s) { if (s == (int)lane) my_t = t; t += calc_dt(t, c); }
	t_base = t;
	return my_t;
}

constexpr uint32_t kMaskWindows = 64;   // windows per ray whose emit masks the count pass keeps for the emit pass

// Marches one ray with the whole warp; returns the number of samples (<= n_limit).  EMIT writes them to out[0..n).
// RECORD keeps, for the first kMaskWindows windows, which lattice points emitted (lane w holds window w / window 32+w) and the
// number of windows the march took, so the emit pass can place the samples without touching the occupancy grid again.
template <bool EMIT, bool RECORD>
__device__ __forceinline__ uint32_t march_ray(const ngp_sampler_cfg& c, const RaySetup& r, const uint8_t* __restrict__ bits, uint32_t lane, uint32_t n_limit,
                                              Coordinate* __restrict__ out, uint32_t& rec_lo, uint32_t& rec_hi, uint32_t& n_windows) {
	float t_base = r.startt;
	float skip_target = -INFINITY;       // lattice points below it are the ones the reference's do-while steps over
	uint32_t j = 0;
	if (RECORD) { rec_lo = rec_hi = 0; n_windows = 0; }
	if (n_limit == 0) return 0;
	const bool const_dt = c.cone_angle == 0.f;
	for (uint32_t window = 0; window < (1u << 16); ++window) {      // bound: no ray of a sane camera needs 2 M lattice points
		const float my_t = lattice_window(t_base, lane, c, const_dt);
		const V3 pos = ray_at(r.o, r.d, my_t);
		const bool inside = aabb_contains(c, pos);
		const float dt = calc_dt(my_t, c);
		bool occ = false;
		float t_target = 0.f;
		if (inside) {
			const uint32_t mip = mip_from_dt(dt, pos, c);
			occ = grid_occupied(pos, bits, mip, c.grid_size);
			if (!occ) t_target = my_t + distance_to_next_voxel(pos, r.d, r.idir, c.grid_size >> mip);
		}
		const uint32_t inside_mask = __ballot_sync(0xffffffffu, inside);
		const uint32_t run_mask = __ballot_sync(0xffffffffu, inside && occ);     // points that emit when visited
		uint32_t p = 0, emitted = 0;
		bool done = false;
		while (p < 32) {
			const uint32_t cand = __ballot_sync(0xffffffffu, lane >= p && !(my_t < skip_target));
			if (!cand) break;                                      // rest of the window is skipped; the target carries over
			const uint32_t q = (uint32_t)__ffs(cand) - 1u;
			if (!((inside_mask >> q) & 1u)) { done = true; break; } // the ray left the box: while (contains(pos) ...) ends
			if ((run_mask >> q) & 1u) {
				const uint32_t rest = ~run_mask >> q;              // first later point that does not emit (bit 0 is clear)
				const uint32_t run = rest ? (uint32_t)__ffs(rest) - 1u : 32u - q;
				const uint32_t take = min(run, n_limit - j);
				if (EMIT && lane >= q && lane < q + take) out[j + (lane - q)] = make_coord(pos, r.d, dt, c);
				if (RECORD) emitted |= (take >= 32u ? 0xffffffffu : ((1u << take) - 1u)) << q;
				j += take;
				if (j >= n_limit) { done = true; break; }
				p = q + run;
				skip_target = -INFINITY;                           // after a sample the very next lattice point is visited
			} else {
				skip_target = __shfl_sync(0xffffffffu, t_target, q);
				p = q + 1;                                         // the do-while advances at least once
			}
		}
		if (RECORD) {
			if (window < 32) { if (lane == window) rec_lo = emitted; }
			else if (window < kMaskWindows) { if (lane == window - 32) rec_hi = emitted; }
			n_windows = window + 1;
		}
		if (done) return j;
	}
	return j;
}

// Emit pass driven by the recorded masks: the lattice is rebuilt (it does not depend on the grid) and every emitting point
// writes its coordinate at the rank of its bit; no bitfield loads, no resolve.
__device__ __forceinline__ void emit_from_masks(const ngp_sampler_cfg& c, const RaySetup& r, uint32_t lane, uint32_t n_windows, uint32_t rec_lo, uint32_t rec_hi,
                                                Coordinate* __restrict__ out) {
	float t_base = r.startt;
	uint32_t j = 0;
	const bool const_dt = c.cone_angle == 0.f;
	const uint32_t lt = (1u << lane) - 1u;
	for (uint32_t window = 0; window < n_windows; ++window) {
		const float my_t = lattice_window(t_base, lane, c, const_dt);
		const uint32_t mask = __shfl_sync(0xffffffffu, window < 32 ? rec_lo : rec_hi, window & 31);
		if (!mask) continue;
		if ((mask >> lane) & 1u) out[j + __popc(mask & lt)] = make_coord(ray_at(r.o, r.d, my_t), r.d, calc_dt(my_t, c), c);
		j += __popc(mask);
	}
}

__global__ void __launch_bounds__(256) march_count_kernel(
	ngp_sampler_cfg c, ngp_camera_cfg cam, uint32_t n_rays, uint32_t n_rays_shift, uint32_t n_images, const float* __restrict__ xforms,
	const uint8_t* __restrict__ bits, int snap, Pcg32 rng, uint32_t* __restrict__ n_out, uint32_t* __restrict__ per_ray_n,
	uint32_t* __restrict__ n_windows_out, uint32_t* __restrict__ masks_out)
{
	const uint32_t lane = threadIdx.x & 31, warps = gridDim.x * (blockDim.x >> 5);
	for (uint32_t i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n_rays; i += warps) {
		const RaySetup r = setup_ray(c, cam, i, n_rays, n_rays_shift, n_images, xforms, snap, rng);
		uint32_t rec_lo, rec_hi, nw;
		const uint32_t n = march_ray<false, true>(c, r, bits, lane, c.max_steps, nullptr, rec_lo, rec_hi, nw);
		if (lane == 0) { n_out[i] = n; n_windows_out[i] = nw; if (per_ray_n) per_ray_n[i] = n; }
		if (n > 0 && nw <= kMaskWindows) {
			if (lane < nw) masks_out[(size_t)i * kMaskWindows + lane] = rec_lo;
			if (lane + 32 < nw) masks_out[(size_t)i * kMaskWindows + 32 + lane] = rec_hi;
		}
	}
}

// exclusive scan in ray order of (ray has samples, sample count); one CTA.  counters[0] = rays kept, counters[1] = samples
// counted including dropped rays (what the reference's atomicAdd would have reached).
__global__ void __launch_bounds__(1024) march_scan_kernel(uint32_t n_rays, uint32_t n_coords, const uint32_t* __restrict__ n_in, uint32_t* __restrict__ base_out,
                                                          uint32_t* __restrict__ ridx_out, uint32_t* __restrict__ counters, uint32_t keep_empty) {
	__shared__ unsigned long long warp_tot[32];
	__shared__ uint32_t kept_total;
	if (threadIdx.x == 0) kept_total = 0;
	const uint32_t per = (n_rays + 1023) / 1024, lo = min(threadIdx.x * per, n_rays), hi = min(lo + per, n_rays);
	unsigned long long sum = 0;
	// keep_empty (train_envmap, occupancy_sampler.cu:408): a ray without samples still takes a slot in the ray list
	for (uint32_t j = lo; j < hi; ++j) { const uint32_t n = n_in[j]; sum += ((unsigned long long)(n > 0 || keep_empty) << 32) | n; }
	const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	unsigned long long incl = sum;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) { const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
	if (lane == 31) warp_tot[warp] = incl;
	__syncthreads();
	if (warp == 0) {
		const unsigned long long w = warp_tot[lane];
		unsigned long long wi = w;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const unsigned long long t = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += t; }
		warp_tot[lane] = wi - w;
		if (lane == 31) counters[1] = (uint32_t)wi;
	}
	__syncthreads();
	unsigned long long run = warp_tot[warp] + incl - sum;
	uint32_t kept = 0;
	for (uint32_t j = lo; j < hi; ++j) {
		const uint32_t n = n_in[j], base = (uint32_t)run;
		base_out[j] = base; ridx_out[j] = (uint32_t)(run >> 32);
		if ((n > 0 || keep_empty) && (uint64_t)base + n <= n_coords) ++kept;
		run += ((unsigned long long)(n > 0 || keep_empty) << 32) | n;
	}
	if (kept) atomicAdd(&kept_total, kept);
	__syncthreads();
	if (threadIdx.x == 0) counters[0] = kept_total;
}

__global__ void __launch_bounds__(256) march_emit_kernel(
	ngp_sampler_cfg c, ngp_camera_cfg cam, uint32_t n_rays, uint32_t n_rays_shift, uint32_t n_coords, uint32_t n_images,
	const float* __restrict__ xforms, const uint2* __restrict__ images, const uint8_t* __restrict__ bits, int snap, Pcg32 rng,
	const uint32_t* __restrict__ n_in, const uint32_t* __restrict__ base_in, const uint32_t* __restrict__ ridx_in,
	const uint32_t* __restrict__ n_windows_in, const uint32_t* __restrict__ masks_in, float* __restrict__ rays_out, float4* __restrict__ rgba_out, uint32_t* __restrict__ numsteps_out, Coordinate* __restrict__ coords_out)
{
	const bool keep_empty = c.keep_empty_rays != 0;
	const uint32_t lane = threadIdx.x & 31, warps = gridDim.x * (blockDim.x >> 5);
	for (uint32_t i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); i < n_rays; i += warps) {
	const uint32_t n = n_in[i], base = base_in[i], ray_idx = ridx_in[i];
	if ((n == 0 && !keep_empty) || (uint64_t)base + n > n_coords) continue;    // no samples, or over budget: dropped like the reference does
	const RaySetup r = setup_ray(c, cam, i, n_rays, n_rays_shift, n_images, xforms, snap, rng);
	if (lane == 0) {
		float* ro = rays_out + (size_t)ray_idx * 6;
		ro[0] = r.o.x; ro[1] = r.o.y; ro[2] = r.o.z; ro[3] = r.d.x; ro[4] = r.d.y; ro[5] = r.d.z;
		const int px = max(min((int)(r.u * (float)cam.width), cam.width - 1), 0), py = max(min((int)(r.v * (float)cam.height), cam.height - 1), 0);
		const uint2 pix = __ldg(images + ((size_t)px + (size_t)py * cam.width + (size_t)r.img * cam.width * cam.height));
		const float2 rg = __half22float2(*reinterpret_cast<const __half2*>(&pix.x)), ba = __half22float2(*reinterpret_cast<const __half2*>(&pix.y));
		rgba_out[ray_idx] = make_float4(rg.x, rg.y, ba.x, ba.y);
		numsteps_out[(size_t)ray_idx * 2 + 0] = n;
		numsteps_out[(size_t)ray_idx * 2 + 1] = base;
	}
	if (n == 0) continue;                                   // kept for the environment map only: a ray record, no samples
	const uint32_t nw = n_windows_in[i];
	if (nw <= kMaskWindows) {
		const uint32_t rec_lo = lane < nw ? masks_in[(size_t)i * kMaskWindows + lane] : 0u;
		const uint32_t rec_hi = lane + 32 < nw ? masks_in[(size_t)i * kMaskWindows + 32 + lane] : 0u;
		emit_from_masks(c, r, lane, nw, rec_lo, rec_hi, coords_out + base);
	} else {
		uint32_t a, b, w;
		march_ray<true, false>(c, r, bits, lane, n, coords_out + base, a, b, w);   // very long rays: march again
	}
	}
}

} // namespace ngp

using namespace ngp;

// scratch of ngp_generate_training_samples: per ray (count, base, compacted ray index, window count, emit masks)
extern "C" size_t ngp_scan_workspace_bytes(uint32_t n) { return (size_t)n * (16 + kMaskWindows * 4) + 256; }

extern "C" int ngp_generate_training_samples(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, uint32_t n_rays, uint32_t n_rays_shift,
                                             uint32_t n_coords, uint32_t n_images, const float* d_xforms, const void* d_images, const uint8_t* d_bitfield,
                                             int snap_to_pixel_centers, uint64_t rng_state, uint64_t rng_inc, float* d_rays, float* d_rays_rgba,
                                             uint32_t* d_numsteps, float* d_coords, uint32_t* d_counters, uint32_t* d_per_ray_n, void* d_scan_ws) {
	NGP_REQUIRE(cfg && cam && d_xforms && d_images && d_bitfield && d_rays && d_rays_rgba && d_numsteps && d_coords && d_counters && d_scan_ws, "null pointer");
	NGP_REQUIRE(n_rays > 0 && n_images > 0, "n_rays and n_images must be positive");
	cudaStream_t s = (cudaStream_t)stream;
	Pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
	uint32_t* n_tmp = (uint32_t*)d_scan_ws;
	uint32_t* base_tmp = n_tmp + n_rays;
	uint32_t* ridx_tmp = base_tmp + n_rays;
	uint32_t* nwin_tmp = ridx_tmp + n_rays;
	uint32_t* mask_tmp = nwin_tmp + n_rays;
	// at most 4 CTAs (32 warps) per SM: the march is latency-bound and usually runs underneath another stream's kernels
	const uint32_t ctas_per_sm = 4u;
	const uint32_t grid = std::min(div_up(n_rays, 8), kNumSMs * ctas_per_sm);
	NGP_LAUNCH("march_count_kernel", s, march_count_kernel<<<grid, 256, 0, s>>>(*cfg, *cam, n_rays, n_rays_shift, n_images, d_xforms, d_bitfield,
		snap_to_pixel_centers, rng, n_tmp, d_per_ray_n, nwin_tmp, mask_tmp));
	NGP_LAUNCH("march_scan_kernel", s, march_scan_kernel<<<1, 1024, 0, s>>>(n_rays, n_coords, n_tmp, base_tmp, ridx_tmp, d_counters, cfg->keep_empty_rays));
	NGP_LAUNCH("march_emit_kernel", s, march_emit_kernel<<<grid, 256, 0, s>>>(*cfg, *cam, n_rays, n_rays_shift, n_coords, n_images, d_xforms, (const uint2*)d_images,
		d_bitfield, snap_to_pixel_centers, rng, n_tmp, base_tmp, ridx_tmp, nwin_tmp, mask_tmp, d_rays, (float4*)d_rays_rgba, d_numsteps, (Coordinate*)d_coords));
	return check_launch("generate_training_samples");
}
