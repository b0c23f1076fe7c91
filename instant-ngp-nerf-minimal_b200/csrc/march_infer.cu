// Inference-side sampler kernels for sm_100a: camera ray set-up, skip to the first occupied cell, per-iteration sample
// generation.  Replaces init_rays_with_payload_kernel_nerf (occupancy_sampler.cu:451), advance_pos_nerf (:505) -- fused
// into one kernel -- and generate_coords_from_rays_regularly_kernel (:552).  Live-ray counts are read from device
// memory so the render loop needs no host round trip per iteration (the reference syncs every iteration, testbed.cu:540).
#include "march.cuh"
#include "composite.cuh"

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

// one 8-point window of a ray's stepping lattice for an octet of lanes (see generate_coords_from_rays_kernel below)
__device__ __forceinline__ float octet_lattice(float t_base, uint32_t sub, uint32_t shift, uint32_t gmask, const ngp_sampler_cfg& c, bool const_dt, float& t_next_base) {
	if (const_dt) {
		const float dt0 = calc_dt(t_base, c);
		const float dtq = (t_base + dt0) - t_base;
		const float cand = __fmaf_rn((float)sub, dtq, t_base), next = __fmaf_rn((float)(sub + 1), dtq, t_base);
		if ((__ballot_sync(gmask, cand + dt0 == next) & gmask) == gmask) { t_next_base = __shfl_sync(gmask, next, shift + 7); return cand; }
	}
	float t = t_base, my_t = t_base;
	#pragma unroll
	for (int s = 0; s < 8; ++s) { if (s == (int)sub) my_t = t; t += calc_dt(t, c); }
	t_next_base = t;
	return my_t;
}

// One thread per pixel: ray set-up, then the skip to the first occupied cell (advance_pos_nerf).  An 8-lanes-per-pixel variant of
// this kernel was measured 6x slower (1.2 -> 7.1 ms at 1080p): the skip crosses about one voxel per 8-point window, so the
// rounds do not shrink while the thread count grows 8x.
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
	float dt = calc_dt(t, c);
	t = __fmaf_rn(ld_random_val(spp, idx * 786433u, 0), dt, t);
	V3 pos;
	while (true) {
		if (!aabb_contains(c, pos = ray_at(o, d, t))) { p.alive = 0; break; }
		dt = calc_dt(t, c);
		const uint32_t mip = mip_from_dt(dt, pos, c);
		if (grid_occupied(pos, bits, mip, c.grid_size)) break;
		t = advance_to_next_voxel(t, c, pos, d, idir, c.grid_size >> mip);
	}
	p.t = t;
	payloads[slot] = p;
}

// 8 lanes per ray ("octet"): the ray's stepping lattice (see march_train.cu) is taken 8 points at a time, every lane evaluates
// one point (position, cascade, occupancy bit, exit distance of its voxel), and an octet-local ballot replays the reference's
// sequential control flow: occupied visited points become samples, an empty visited point skips to the first lattice point
// behind its voxel.  Sample positions, step counts and the stored ray parameter are bit-identical to the thread-per-ray loop,
// but the 8 Coordinates of a ray leave as one contiguous 224-byte store, four rays share a warp instead of 32 (the reference's
// loop lengths differ wildly from ray to ray), and empty space is crossed 8 lattice points per round.
__global__ void __launch_bounds__(256) generate_coords_from_rays_kernel(ngp_sampler_cfg c, uint32_t n_rays_max, const uint32_t* __restrict__ n_rays_ptr, uint32_t n_steps,
                                                                         const uint8_t* __restrict__ bits, RayPayload* __restrict__ payloads, Coordinate* __restrict__ coords) {
	const uint32_t n_rays = n_rays_ptr ? min(*n_rays_ptr, n_rays_max) : n_rays_max;
	const uint32_t lane = threadIdx.x & 31, sub = lane & 7, shift = lane & 24;
	const uint32_t gmask = 0xffu << shift;
	const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 3;
	if (i >= n_rays) return;                                   // the whole octet leaves together
	RayPayload& p = payloads[i];
	if (!p.alive) return;
	const V3 o = {p.ox, p.oy, p.oz}, d = {p.dx, p.dy, p.dz}, idir = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
	float t_base = p.t, skip_target = -INFINITY;
	uint32_t j = 0;
	if (n_steps == 0) { if (sub == 0) p.n_steps = 0; return; }
	const bool const_dt = c.cone_angle == 0.f;
	for (uint32_t window = 0; window < (1u << 18); ++window) {
		float t_next_base;
		const float my_t = octet_lattice(t_base, sub, shift, gmask, c, const_dt, t_next_base);
		const V3 pos = ray_at(o, d, my_t);
		const bool inside = aabb_contains(c, pos);
		const float dt = calc_dt(my_t, c);
		bool occ = false;
		float t_target = 0.f;
		if (inside) {
			const uint32_t mip = mip_from_dt(dt, pos, c);
			occ = grid_occupied(pos, bits, mip, c.grid_size);
			if (!occ) t_target = my_t + distance_to_next_voxel(pos, d, idir, c.grid_size >> mip);
		}
		const uint32_t inside_mask = (__ballot_sync(gmask, inside) >> shift) & 0xffu;
		const uint32_t run_mask = (__ballot_sync(gmask, inside && occ) >> shift) & 0xffu;
		uint32_t q0 = 0;
		while (q0 < 8) {
			const uint32_t cand = (__ballot_sync(gmask, sub >= q0 && !(my_t < skip_target)) >> shift) & 0xffu;
			if (!cand) break;                                      // the rest of the window is stepped over; the target carries on
			const uint32_t q = (uint32_t)__ffs(cand) - 1u;
			if (!((inside_mask >> q) & 1u)) {                      // the ray left the box: n_steps = samples so far, t stays (:581-584)
				if (sub == 0) p.n_steps = j;
				return;
			}
			if ((run_mask >> q) & 1u) {
				const uint32_t rest = (~run_mask & 0xffu) >> q;    // bit 0 is clear
				const uint32_t run = rest ? (uint32_t)__ffs(rest) - 1u : 8u - q;
				const uint32_t take = min(run, n_steps - j);
				if (sub >= q && sub < q + take) coords[(size_t)i * n_steps + j + (sub - q)] = make_coord(pos, d, dt, c);
				j += take;
				if (j >= n_steps) {                                // t after the last sample is the next lattice point
					const float t_after = __shfl_sync(gmask, my_t, shift + ((q + take) & 7u));
					if (sub == 0) { p.t = q + take < 8 ? t_after : t_next_base; p.n_steps = n_steps; }
					return;
				}
				q0 = q + run;
				skip_target = -INFINITY;
			} else {
				skip_target = __shfl_sync(gmask, t_target, shift + q);
				q0 = q + 1;
			}
		}
		t_base = t_next_base;
	}
	if (sub == 0) { p.t = t_base; p.n_steps = j; }                 // unreachable for sane cameras (2 M lattice points)
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

extern "C" int ngp_generate_coords_from_rays(void* stream, const ngp_sampler_cfg* cfg, uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps,
                                             const uint8_t* d_bitfield, void* d_payloads, float* d_coords) {
	NGP_REQUIRE(cfg && d_bitfield && d_payloads && d_coords, "null pointer");
	if (n_rays == 0) return 0;
	NGP_LAUNCH("generate_coords_from_rays_kernel", (cudaStream_t)stream, generate_coords_from_rays_kernel<<<div_up((uint64_t)n_rays * 8, 256), 256, 0, (cudaStream_t)stream>>>(*cfg, n_rays, d_n_rays, n_steps, d_bitfield, (RayPayload*)d_payloads, (Coordinate*)d_coords));
	return check_launch("generate_coords_from_rays");
}
