// Inference-side sampler kernels for sm_100a: camera ray set-up, skip to the first occupied cell, per-iteration sample
// generation.  Replaces init_rays_with_payload_kernel_nerf (occupancy_sampler.cu:451), advance_pos_nerf (:505) -- fused
// into one kernel -- and generate_coords_from_rays_regularly_kernel (:552).  Live-ray counts are read from device
// memory so the render loop needs no host round trip per iteration (the reference syncs every iteration, testbed.cu:540).
#include "march.cuh"
#include "composite.cuh"
#include <algorithm>
#include <atomic>

namespace ngp {

// Sobol' dimensions 0 and 1 + Laine-Karras scrambling (internal/utils/random_val.cuh:142-278, after Burley 2020)
__host__ __device__ inline uint32_t sobol01(uint32_t index, uint32_t dim) {
	uint32_t X = 0, v = 0x80000000u;
	for (uint32_t bit = 0; bit < 32; ++bit) {
		const uint32_t dirn = dim == 0 ? (0x80000000u >> bit) : v;
		if ((index >> bit) & 1) X ^= dirn;
		v = v ^ (v >> 1);
	}
	return X;
}
__host__ __device__ inline uint32_t hash_combine(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
__host__ __device__ inline uint32_t reverse_bits32(uint32_t x) {
	x = ((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1);
	x = ((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2);
	x = ((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4);
	x = ((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8);
	return (x >> 16) | (x << 16);
}
__host__ __device__ inline uint32_t laine_karras(uint32_t x, uint32_t seed) {
	x += seed; x ^= x * 0x6c50b47cu; x ^= x * 0xb82f1e52u; x ^= x * 0xc7afe638u; x ^= x * 0x8d22f6e6u;
	return x;
}
__host__ __device__ inline uint32_t nus_base2(uint32_t x, uint32_t seed) { return reverse_bits32(laine_karras(reverse_bits32(x), seed)); }
__host__ __device__ inline float ld_random_val(uint32_t index, uint32_t seed, uint32_t dim) {
	const float S = 2.3283064365386963e-10f; // 2^-32
	index = nus_base2(index, seed);
	return (float)nus_base2(sobol01(index, dim), hash_combine(seed, dim)) * S;
}

struct CamMat { float m[12]; };

// The reference's per-ray loop (generate_coords_from_rays_regularly_kernel, occupancy_sampler.cu:552-601) on persistent warps that
// REFILL: a lane whose ray is finished takes the next unprocessed ray, so a warp is not held up by its longest ray.
//   * ncu on the thread-per-ray form of this loop (profiles/r02_render.md): 8.1 of 32 lanes active -- rays of one warp need between 4
//     and a few hundred loop iterations per round (empty voxels to cross between samples) and the short ones idle.
//   * rays are dealt out dynamically: a warp draws blocks of 32 ray indices from one device-wide counter (one atomic per 32 rays)
//     and hands them to its idle lanes in index order; the last warp to leave resets the counter for the next launch.
//   * one loop iteration = one visited lattice point: ray_at, cascade, occupancy bit, then either a sample or the empty-voxel hop in
//     closed form (lattice_hop.h).  make_coord's per-ray parts are hoisted; a Coordinate leaves as 8-byte stores where aligned.
// Sample positions, step counts and the stored ray parameter are bit-identical to the reference loop (the visiting order of a ray's
// lattice points is the reference's; only which lane works on which ray differs).
constexpr uint32_t kGenQueues = 64;
__device__ uint32_t g_gen_queue[kGenQueues * 2];       // per launch slot: [0] next ray index, [1] warps that have left

struct MarchRay {
	V3 o, d, idir, half_sign;          // half_sign = 0.5 * copysign(1, d): distance_to_next_voxel's per-axis constant
	float cdx, cdy, cdz, t;
	ngp_lattice lat;                   // closed-form stepping constants of the binade t is in (lattice_hop.h)
	uint32_t index, j, known_cell;
};

// a Coordinate (28 bytes, 4-byte aligned) as 8-byte stores where the alignment allows
__device__ __forceinline__ void store_coordinate(Coordinate* dst, float px, float py, float pz, float dt, float dx, float dy, float dz) {
	float* f = reinterpret_cast<float*>(dst);
	if ((reinterpret_cast<uintptr_t>(f) & 7u) == 0) {
		reinterpret_cast<float2*>(f)[0] = make_float2(px, py); reinterpret_cast<float2*>(f)[1] = make_float2(pz, dt);
		reinterpret_cast<float2*>(f)[2] = make_float2(dx, dy); f[6] = dz;
	} else {
		f[0] = px; reinterpret_cast<float2*>(f + 1)[0] = make_float2(py, pz); reinterpret_cast<float2*>(f + 1)[1] = make_float2(dt, dx);
		reinterpret_cast<float2*>(f + 1)[2] = make_float2(dy, dz);
	}
}

__global__ void __launch_bounds__(256) generate_coords_from_rays_kernel(ngp_sampler_cfg c, uint32_t n_rays_max, const uint32_t* __restrict__ n_rays_ptr,
                                                                         uint32_t n_steps_arg, const uint32_t* __restrict__ n_steps_ptr, uint32_t queue_slot,
                                                                         const uint8_t* __restrict__ bits, RayPayload* __restrict__ payloads, Coordinate* __restrict__ coords) {
	__shared__ uint32_t spread_table[256];
	const uint32_t* spread = nullptr;
	if (c.grid_size <= 256) {
		spread_table[threadIdx.x] = morton3D(threadIdx.x, 0, 0);
		__syncthreads();
		spread = spread_table;
	}
	const uint32_t n_rays = n_rays_ptr ? min(*n_rays_ptr, n_rays_max) : n_rays_max;
	const uint32_t n_steps = n_steps_ptr ? *n_steps_ptr : n_steps_arg;
	const uint32_t lane = threadIdx.x & 31, lt_mask = (1u << lane) - 1u;
	uint32_t* queue = g_gen_queue + 2 * queue_slot;
	const bool const_dt = c.cone_angle == 0.f;
	const float diag = c.aabb_max - c.aabb_min;
	int diag_exp;
	const bool diag_pow2 = frexpf(diag, &diag_exp) == 0.5f;      // then x / diag == x * (1 / diag) exactly
	const float inv_diag = 1.0f / diag;
	float wdt_const = warp_dt(c.min_step, c);
	asm volatile("" : "+f"(wdt_const));                            // keep the division out of the loop (nvcc folds it back into the sample path otherwise)
	const float coord_ulp = ngp_binade_top(fmaxf(fmaxf(fabsf(c.aabb_min), fabsf(c.aabb_max)), 1.f)) * 5.9604645e-8f;

	MarchRay r;
	bool have = false, drained = false;
	uint32_t wnext = 0, wend = 0;                                  // this warp's private block of ray indices (warp-uniform)
	while (true) {
		// Refill in batches: the set-up of a ray (payload load, three reciprocals) runs with only the idle lanes active, so it waits
		// until a quarter of the warp is idle -- or nothing is left to wait for
		const uint32_t need = __ballot_sync(0xffffffffu, !have);
		if (__popc(need) >= 8 || (need && drained)) {
			if (wnext >= wend && !drained) {
				uint32_t b = 0;
				if (lane == 0) b = atomicAdd(queue, 32u);
				b = __shfl_sync(0xffffffffu, b, 0);
				if (b >= n_rays) { drained = true; wnext = wend = 0; }
				else { wnext = b; wend = min(b + 32u, n_rays); }
			}
			if (wnext < wend) {
				const uint32_t cand = wnext + __popc(need & lt_mask);
				wnext += __popc(need);
				if (!have && cand < wend) {
					const uint2* src = reinterpret_cast<const uint2*>(payloads + cand);      // 40-byte records, 8-byte aligned
					const uint2 w0 = src[0], w1 = src[1], w2 = src[2], w3 = src[3], w4 = src[4];
					if (w0.x) {                                    // alive
						if (n_steps == 0) payloads[cand].n_steps = 0;
						else {
							r.o = {__uint_as_float(w0.y), __uint_as_float(w1.x), __uint_as_float(w1.y)};
							r.d = {__uint_as_float(w2.x), __uint_as_float(w2.y), __uint_as_float(w3.x)};
							r.idir = {1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z};
							r.half_sign = {0.5f * copysignf(1.f, r.d.x), 0.5f * copysignf(1.f, r.d.y), 0.5f * copysignf(1.f, r.d.z)};
							r.cdx = (r.d.x + 1.f) * 0.5f; r.cdy = (r.d.y + 1.f) * 0.5f; r.cdz = (r.d.z + 1.f) * 0.5f;
							r.t = __uint_as_float(w4.x); r.index = cand; r.j = 0; r.known_cell = 0xffffffffu;
							ngp_lattice_enter(&r.lat, r.t, c.min_step, const_dt);
							have = true;
						}
					}
				}
			} else if (drained && need == 0xffffffffu) break;
		}
		if (have) {
			const V3 pos = ray_at(r.o, r.d, r.t);
			if (!aabb_contains(c, pos)) {                          // the ray left the box: n_steps = samples so far, t stays (:581-584)
				payloads[r.index].n_steps = r.j;
				have = false;
			} else {
				const float dt = calc_dt(r.t, c);
				const uint32_t mip = mip_at(dt, pos, c);
				const CellProbe cp = probe_cell(pos, mip, c.grid_size, spread, bits, &r.known_cell);
				if (cp.occupied) {
					float px, py, pz;
					if (diag_pow2) { px = (pos.x - c.aabb_min) * inv_diag; py = (pos.y - c.aabb_min) * inv_diag; pz = (pos.z - c.aabb_min) * inv_diag; }
					else { px = (pos.x - c.aabb_min) / diag; py = (pos.y - c.aabb_min) / diag; pz = (pos.z - c.aabb_min) / diag; }
					float wdt = wdt_const;
					if (!const_dt) wdt = warp_dt(dt, c);
					store_coordinate(coords + (size_t)r.index * n_steps + r.j, px, py, pz, wdt, r.cdx, r.cdy, r.cdz);
					++r.j;
					r.t += dt;
					if (r.j >= n_steps) {
						payloads[r.index].t = r.t; payloads[r.index].n_steps = n_steps;
						have = false;
					} else if (!(r.t < r.lat.top)) ngp_lattice_enter(&r.lat, r.t, c.min_step, const_dt);
				} else if (!coarse_hop(c, r.lat, r.t, cp, mip, r.d, r.idir, coord_ulp)) {
					cell_hop(c, r.lat, r.t, pos, r.idir, r.half_sign, mip, const_dt);
				}
			}
		}
	}
	// the last warp out re-arms the queue for the next launch that uses this slot
	if (lane == 0) {
		__threadfence();
		const uint32_t total_warps = gridDim.x * (blockDim.x >> 5);
		if (atomicAdd(queue + 1, 1u) == total_warps - 1u) { queue[0] = 0; queue[1] = 0; __threadfence(); }
	}
}

// One thread per pixel: ray set-up, then the skip to the first occupied cell (advance_pos_nerf) with the closed-form hop of
// lattice_hop.h and whole empty blocks crossed at once (coarse_hop).  An 8-lanes-per-pixel variant of this kernel was measured 6x slower (1.2 -> 7.1 ms at 1080p): the skip crosses
// about one voxel per 8-point window, so the rounds do not shrink while the thread count grows 8x.
__global__ void __launch_bounds__(128) init_rays_kernel(ngp_sampler_cfg c, ngp_camera_cfg cam, CamMat cam_matrix, float offx, float offy,
                                                        uint32_t spp, const uint8_t* __restrict__ bits, RayPayload* __restrict__ payloads,
                                                        uint32_t first_pixel, uint32_t n, const __half* __restrict__ envmap, int env_res_x, int env_res_y,
                                                        float4* __restrict__ frame) {
	// tile-sharded rendering: the launch covers the n pixels from first_pixel on (whole rows); every quantity that depends on the
	// pixel -- the ray, the jitter seed, the scatter target -- uses the GLOBAL pixel index, so a band is bit-identical to the same
	// rows of the full frame; only the slot in the payload array is band-relative
	const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
	if (slot >= n) return;
	const uint32_t idx = first_pixel + slot;
	const uint32_t x = idx % cam.width, y = idx / cam.width;
	V3 o, d;
	make_ray(cam_matrix.m, ((float)x + offx) / (float)cam.width, ((float)y + offy) / (float)cam.height, cam, o, d);
	if (envmap) {            // occupancy_sampler.cu:500-502: every pixel starts from the environment map seen along its ray
		float ev[4];
		read_envmap(envmap, env_res_x, env_res_y, d.x, d.y, d.z, ev);
		frame[idx] = make_float4(ev[0], ev[1], ev[2], ev[3]);
	}
	float t = fmaxf(ray_intersect_tmin(c, o, d), 0.05f) + 1e-6f;
	RayPayload p;
	p.ox = o.x; p.oy = o.y; p.oz = o.z; p.dx = d.x; p.dy = d.y; p.dz = d.z; p.idx = idx; p.n_steps = 0; p.t = t; p.alive = 1;
	if (!aabb_contains(c, ray_at(o, d, t))) {
		p.alive = 0; p.dx = p.dy = p.dz = 0.f; p.t = 0.f; p.idx = 0;
		payloads[slot] = p;
		return;
	}
	const V3 idir = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
	const V3 half_sign = {0.5f * copysignf(1.f, d.x), 0.5f * copysignf(1.f, d.y), 0.5f * copysignf(1.f, d.z)};
	const bool const_dt = c.cone_angle == 0.f;
	const float coord_ulp = ngp_binade_top(fmaxf(fmaxf(fabsf(c.aabb_min), fabsf(c.aabb_max)), 1.f)) * 5.9604645e-8f;
	float dt = calc_dt(t, c);
	t = __fmaf_rn(ld_random_val(spp, idx * 786433u, 0), dt, t);
	ngp_lattice lat;
	ngp_lattice_enter(&lat, t, c.min_step, const_dt);
	while (true) {
		const V3 pos = ray_at(o, d, t);
		if (!aabb_contains(c, pos)) { p.alive = 0; break; }
		dt = calc_dt(t, c);
		const uint32_t mip = mip_at(dt, pos, c);
		const CellProbe cp = probe_cell(pos, mip, c.grid_size, nullptr, bits);
		if (cp.occupied) break;
		if (!coarse_hop(c, lat, t, cp, mip, d, idir, coord_ulp)) cell_hop(c, lat, t, pos, idir, half_sign, mip, const_dt);
	}
	p.t = t;
	payloads[slot] = p;
}

} // namespace ngp

using namespace ngp;

static float fractf_host(float x) { return x - floorf(x); }

static int init_rays_impl(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, const float* camera_matrix_host, uint32_t spp,
                          int snap_to_pixel_centers, const uint8_t* d_bitfield, void* d_payloads, uint32_t row_begin, uint32_t row_end,
                          const void* d_envmap, int env_res_x, int env_res_y, float* d_frame_buffer) {
	NGP_REQUIRE(cfg && cam && camera_matrix_host && d_bitfield && d_payloads, "null pointer");
	NGP_REQUIRE(row_begin <= row_end && row_end <= (uint32_t)cam->height, "row range outside the image");
	if (row_begin == row_end) return 0;
	// ld_random_pixel_offset (random_val.cuh:273-278): identical for every pixel, so it is evaluated once on the host
	const uint32_t s1 = snap_to_pixel_centers ? 0u : spp;
	const float ax = ld_random_val(0, 0xdeadbeefu, 0), ay = ld_random_val(0, 0xdeadbeefu, 1);
	const float bx = ld_random_val(s1, 0xdeadbeefu, 0), by = ld_random_val(s1, 0xdeadbeefu, 1);
	const float offx = fractf_host(0.5f - ax + bx), offy = fractf_host(0.5f - ay + by);
	const uint32_t n = (uint32_t)cam->width * (row_end - row_begin), first = (uint32_t)cam->width * row_begin;
	CamMat cm; memcpy(cm.m, camera_matrix_host, sizeof(cm.m));
	NGP_LAUNCH("init_rays_kernel", (cudaStream_t)stream, init_rays_kernel<<<div_up(n, 128), 128, 0, (cudaStream_t)stream>>>(*cfg, *cam, cm, offx, offy, spp, d_bitfield, (RayPayload*)d_payloads, first, n,
		(const __half*)d_envmap, env_res_x, env_res_y, (float4*)d_frame_buffer));
	return check_launch("init_rays");
}

extern "C" int ngp_init_rays_rows(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, const float* camera_matrix_host, uint32_t spp,
                                  int snap_to_pixel_centers, const uint8_t* d_bitfield, void* d_payloads, uint32_t row_begin, uint32_t row_end) {
	return init_rays_impl(stream, cfg, cam, camera_matrix_host, spp, snap_to_pixel_centers, d_bitfield, d_payloads, row_begin, row_end, nullptr, 0, 0, nullptr);
}

// the same with a trained environment map: frame_buffer[pixel] <- read_envmap(ray direction) for every pixel of the rows
// (occupancy_sampler.cu:459-502); shade_kernel later composites the hits over it
extern "C" int ngp_init_rays_rows_envmap(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, const float* camera_matrix_host, uint32_t spp,
                                         int snap_to_pixel_centers, const uint8_t* d_bitfield, void* d_payloads, uint32_t row_begin, uint32_t row_end,
                                         const void* d_envmap, int32_t env_res_x, int32_t env_res_y, float* d_frame_buffer) {
	NGP_REQUIRE(d_envmap && d_frame_buffer && env_res_x > 0 && env_res_y > 0, "envmap, its resolution and the frame buffer are required");
	return init_rays_impl(stream, cfg, cam, camera_matrix_host, spp, snap_to_pixel_centers, d_bitfield, d_payloads, row_begin, row_end, d_envmap, env_res_x, env_res_y, d_frame_buffer);
}

extern "C" int ngp_init_rays(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, const float* camera_matrix_host, uint32_t spp,
                             int snap_to_pixel_centers, const uint8_t* d_bitfield, void* d_payloads) {
	NGP_REQUIRE(cam, "null pointer");
	return ngp_init_rays_rows(stream, cfg, cam, camera_matrix_host, spp, snap_to_pixel_centers, d_bitfield, d_payloads, 0, (uint32_t)cam->height);
}

static int generate_coords_impl(void* stream, const ngp_sampler_cfg* cfg, uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps, const uint32_t* d_n_steps,
                                const uint8_t* d_bitfield, void* d_payloads, float* d_coords) {
	NGP_REQUIRE(cfg && d_bitfield && d_payloads && d_coords, "null pointer");
	if (n_rays == 0) return 0;
	// launches take turns on kGenQueues work counters, so launches that overlap on the device (other streams) do not share one
	static std::atomic<uint32_t> turn{0};
	const uint32_t slot = turn.fetch_add(1) % kGenQueues;
	// Persistent warps that refill.  A round is over when its longest ray is (some rays cross tens of empty voxels between samples),
	// so the warps are kept FEW and full: about 8 rays per lane lets the refill even out the rest, while one warp per ray-lane
	// (ncu, profiles/r02_render.md) spent 2/3 of its issue slots on warps holding one or two unfinished rays.
	const uint32_t blocks = (uint32_t)std::min<uint64_t>((uint64_t)kNumSMs * 2, std::max<uint64_t>(div_up(n_rays, 2048), std::min<uint64_t>(div_up(n_rays, 256), kNumSMs)));
	NGP_LAUNCH("generate_coords_from_rays_kernel", (cudaStream_t)stream, generate_coords_from_rays_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(*cfg, n_rays, d_n_rays, n_steps, d_n_steps, slot, d_bitfield, (RayPayload*)d_payloads, (Coordinate*)d_coords));
	return check_launch("generate_coords_from_rays");
}

extern "C" int ngp_generate_coords_from_rays(void* stream, const ngp_sampler_cfg* cfg, uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps,
                                             const uint8_t* d_bitfield, void* d_payloads, float* d_coords) {
	return generate_coords_impl(stream, cfg, n_rays, d_n_rays, n_steps, nullptr, d_bitfield, d_payloads, d_coords);
}

extern "C" int ngp_generate_coords_from_rays_dev(void* stream, const ngp_sampler_cfg* cfg, uint32_t n_rays_bound, const uint32_t* d_ctl, uint32_t round,
                                                 const uint8_t* d_bitfield, void* d_payloads, float* d_coords) {
	NGP_REQUIRE(d_ctl, "null pointer");
	return generate_coords_impl(stream, cfg, n_rays_bound, d_ctl + ((round + 1) & 1), 0, d_ctl + 2, d_bitfield, d_payloads, d_coords);
}
