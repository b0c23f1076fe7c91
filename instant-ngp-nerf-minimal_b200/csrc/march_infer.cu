// Inference-side sampler kernels for sm_100a: camera ray set-up, skip to the first occupied cell, per-iteration sample
// generation.  Replaces init_rays_with_payload_kernel_nerf (occupancy_sampler.cu:451), advance_pos_nerf (:505) -- fused
// into one kernel -- and generate_coords_from_rays_regularly_kernel (:552).  Live-ray counts are read from device
// memory so the render loop needs no host round trip per iteration (the reference syncs every iteration, testbed.cu:540).
#include "march.cuh"

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

__global__ void __launch_bounds__(128) init_rays_kernel(ngp_sampler_cfg c, ngp_camera_cfg cam, CamMat cam_matrix, float offx, float offy,
                                                        uint32_t spp, const uint8_t* __restrict__ bits, RayPayload* __restrict__ payloads) {
	const uint32_t n = (uint32_t)cam.width * (uint32_t)cam.height;
	const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
	if (idx >= n) return;
	const uint32_t x = idx % cam.width, y = idx / cam.width;
	V3 o, d;
	make_ray(cam_matrix.m, ((float)x + offx) / (float)cam.width, ((float)y + offy) / (float)cam.height, cam, o, d);
	float t = fmaxf(ray_intersect_tmin(c, o, d), 0.05f) + 1e-6f;
	RayPayload p;
	p.ox = o.x; p.oy = o.y; p.oz = o.z; p.dx = d.x; p.dy = d.y; p.dz = d.z; p.idx = idx; p.n_steps = 0; p.t = t; p.alive = 1;
	if (!aabb_contains(c, ray_at(o, d, t))) {
		p.alive = 0; p.dx = p.dy = p.dz = 0.f; p.t = 0.f; p.idx = 0;
		payloads[idx] = p;
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
	payloads[idx] = p;
}

__global__ void __launch_bounds__(128) generate_coords_from_rays_kernel(ngp_sampler_cfg c, uint32_t n_rays_max, const uint32_t* __restrict__ n_rays_ptr, uint32_t n_steps,
                                                                         const uint8_t* __restrict__ bits, RayPayload* __restrict__ payloads, Coordinate* __restrict__ coords) {
	const uint32_t n_rays = n_rays_ptr ? min(*n_rays_ptr, n_rays_max) : n_rays_max;
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n_rays) return;
	RayPayload& p = payloads[i];
	if (!p.alive) return;
	const V3 o = {p.ox, p.oy, p.oz}, d = {p.dx, p.dy, p.dz}, idir = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
	float t = p.t;
	for (uint32_t j = 0; j < n_steps; ++j) {
		V3 pos; float dt = 0.f;
		while (true) {
			if (!aabb_contains(c, pos = ray_at(o, d, t))) { p.n_steps = j; return; }
			dt = calc_dt(t, c);
			const uint32_t mip = mip_from_dt(dt, pos, c);
			if (grid_occupied(pos, bits, mip, c.grid_size)) break;
			t = advance_to_next_voxel(t, c, pos, d, idir, c.grid_size >> mip);
		}
		coords[(size_t)i * n_steps + j] = make_coord(pos, d, dt, c);
		t += dt;
	}
	p.t = t;
	p.n_steps = n_steps;
}

} // namespace ngp

using namespace ngp;

static float fractf_host(float x) { return x - floorf(x); }

extern "C" int ngp_init_rays(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, const float* camera_matrix_host, uint32_t spp,
                             int snap_to_pixel_centers, const uint8_t* d_bitfield, void* d_payloads) {
	NGP_REQUIRE(cfg && cam && camera_matrix_host && d_bitfield && d_payloads, "null pointer");
	// ld_random_pixel_offset (random_val.cuh:273-278): identical for every pixel, so it is evaluated once on the host
	const uint32_t s1 = snap_to_pixel_centers ? 0u : spp;
	const float ax = ld_random_val(0, 0xdeadbeefu, 0), ay = ld_random_val(0, 0xdeadbeefu, 1);
	const float bx = ld_random_val(s1, 0xdeadbeefu, 0), by = ld_random_val(s1, 0xdeadbeefu, 1);
	const float offx = fractf_host(0.5f - ax + bx), offy = fractf_host(0.5f - ay + by);
	const uint32_t n = (uint32_t)cam->width * (uint32_t)cam->height;
	CamMat cm; memcpy(cm.m, camera_matrix_host, sizeof(cm.m));
	NGP_LAUNCH("init_rays_kernel", (cudaStream_t)stream, init_rays_kernel<<<div_up(n, 128), 128, 0, (cudaStream_t)stream>>>(*cfg, *cam, cm, offx, offy, spp, d_bitfield, (RayPayload*)d_payloads));
	return check_launch("init_rays");
}

extern "C" int ngp_generate_coords_from_rays(void* stream, const ngp_sampler_cfg* cfg, uint32_t n_rays, const uint32_t* d_n_rays, uint32_t n_steps,
                                             const uint8_t* d_bitfield, void* d_payloads, float* d_coords) {
	NGP_REQUIRE(cfg && d_bitfield && d_payloads && d_coords, "null pointer");
	if (n_rays == 0) return 0;
	NGP_LAUNCH("generate_coords_from_rays_kernel", (cudaStream_t)stream, generate_coords_from_rays_kernel<<<div_up(n_rays, 128), 128, 0, (cudaStream_t)stream>>>(*cfg, n_rays, d_n_rays, n_steps, d_bitfield, (RayPayload*)d_payloads, (Coordinate*)d_coords));
	return check_launch("generate_coords_from_rays");
}
