// Training ray generation + two-pass occupancy march + ordered compaction, for sm_100a.
// Replaces generate_training_samples_nerf (internal/sampler/occupancy_sampler.cu:310-449).
//
// The reference reserves output space with two global atomicAdds per ray, so its sample order changes from run to run.
// Here each CTA scans its rays' (sample count, ray-kept flag) pair with warp shuffles and chains the CTA totals with a
// single-pass decoupled look-back scan: bases come out in ray order, one 64-bit publish per CTA instead of 2 atomics per
// ray, and the result is bit-reproducible (and equal to the sequential oracle).
#include "march.cuh"

namespace ngp {

__global__ void __launch_bounds__(128) generate_training_samples_kernel(
	ngp_sampler_cfg c, ngp_camera_cfg cam, uint32_t n_rays, uint32_t n_rays_shift, uint32_t n_coords, uint32_t n_images,
	const float* __restrict__ xforms, const uint2* __restrict__ images, const uint8_t* __restrict__ bits, int snap, Pcg32 rng0,
	float* __restrict__ rays_out, float4* __restrict__ rgba_out, uint32_t* __restrict__ numsteps_out, Coordinate* __restrict__ coords_out,
	uint32_t* __restrict__ counters, uint32_t* __restrict__ per_ray_n, ScanState scan)
{
	const uint32_t ticket = scan_ticket(scan);
	const uint32_t i = ticket * blockDim.x + threadIdx.x;
	const bool active = i < n_rays;

	V3 o = {0, 0, 0}, d = {0, 0, 1}, idir = {0, 0, 0};
	float startt = 0.f, u = 0.f, v = 0.f;
	uint32_t img = 0, n = 0;
	if (active) {
		img = (((n_rays_shift + i) * n_images) / n_rays) % n_images;
		Pcg32 rng = rng0;
		rng.advance((uint64_t)i * 8);
		u = rng.next_float(); v = rng.next_float();
		if (snap) {
			u = ((float)min(max((int)(u * (float)cam.width), 0), cam.width - 1) + 0.5f) / (float)cam.width;
			v = ((float)min(max((int)(v * (float)cam.height), 0), cam.height - 1) + 0.5f) / (float)cam.height;
		}
		make_ray(xforms + (size_t)img * 12, u, v, cam, o, d);
		const float tmin = fmaxf(ray_intersect_tmin(c, o, d), c.near_distance);
		startt = __fmaf_rn(calc_dt(tmin, c), rng.next_float(), tmin);
		idir = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
		// pass 1: count occupied steps
		float t = startt; V3 pos;
		while (aabb_contains(c, pos = ray_at(o, d, t)) && n < c.max_steps) {
			const float dt = calc_dt(t, c);
			const uint32_t mip = mip_from_dt(dt, pos, c);
			if (grid_occupied(pos, bits, mip, c.grid_size)) { ++n; t += dt; }
			else t = advance_to_next_voxel(t, c, pos, d, idir, c.grid_size >> mip);
		}
		if (per_ray_n) per_ray_n[i] = n;
	}

	// (rays with samples, samples) packed hi/lo; exclusive prefix in ray order
	const unsigned long long packed = ((unsigned long long)(n > 0 ? 1u : 0u) << 32) | n;
	const unsigned long long prefix = chained_exclusive_scan(scan, ticket, packed);
	const uint32_t base = (uint32_t)(prefix & 0xffffffffull), ray_idx = (uint32_t)(prefix >> 32);

	// the last ray publishes the totals (the sample counter includes dropped rays, like the reference's atomicAdd)
	if (active && i == n_rays - 1) { counters[1] = base + n; }
	const bool kept = active && n > 0 && (uint64_t)base + n <= n_coords;
	// rays kept = number of kept rays; kept rays form a prefix of the rays with samples, so the count is ray_idx+1 of the last kept one
	if (kept) atomicMax(&counters[0], ray_idx + 1);
	if (!kept) return;

	float* ro = rays_out + (size_t)ray_idx * 6;
	ro[0] = o.x; ro[1] = o.y; ro[2] = o.z; ro[3] = d.x; ro[4] = d.y; ro[5] = d.z;
	{
		const int px = max(min((int)(u * (float)cam.width), cam.width - 1), 0), py = max(min((int)(v * (float)cam.height), cam.height - 1), 0);
		const uint2 pix = __ldg(images + ((size_t)px + (size_t)py * cam.width + (size_t)img * cam.width * cam.height));
		const float2 rg = __half22float2(*reinterpret_cast<const __half2*>(&pix.x)), ba = __half22float2(*reinterpret_cast<const __half2*>(&pix.y));
		rgba_out[ray_idx] = make_float4(rg.x, rg.y, ba.x, ba.y);
	}
	numsteps_out[(size_t)ray_idx * 2 + 0] = n;
	numsteps_out[(size_t)ray_idx * 2 + 1] = base;

	// pass 2: emit
	Coordinate* out = coords_out + base;
	uint32_t j = 0; float t = startt; V3 pos;
	while (aabb_contains(c, pos = ray_at(o, d, t)) && j < n) {
		const float dt = calc_dt(t, c);
		const uint32_t mip = mip_from_dt(dt, pos, c);
		if (grid_occupied(pos, bits, mip, c.grid_size)) { out[j] = make_coord(pos, d, dt, c); ++j; t += dt; }
		else t = advance_to_next_voxel(t, c, pos, d, idir, c.grid_size >> mip);
	}
}

} // namespace ngp

using namespace ngp;

extern "C" size_t ngp_scan_workspace_bytes(uint32_t n) { return (size_t)(div_up(n, 128) + 2) * sizeof(unsigned long long); }

extern "C" int ngp_generate_training_samples(void* stream, const ngp_sampler_cfg* cfg, const ngp_camera_cfg* cam, uint32_t n_rays, uint32_t n_rays_shift,
                                             uint32_t n_coords, uint32_t n_images, const float* d_xforms, const void* d_images, const uint8_t* d_bitfield,
                                             int snap_to_pixel_centers, uint64_t rng_state, uint64_t rng_inc, float* d_rays, float* d_rays_rgba,
                                             uint32_t* d_numsteps, float* d_coords, uint32_t* d_counters, uint32_t* d_per_ray_n, void* d_scan_ws) {
	NGP_REQUIRE(cfg && cam && d_xforms && d_images && d_bitfield && d_rays && d_rays_rgba && d_numsteps && d_coords && d_counters && d_scan_ws, "null pointer");
	NGP_REQUIRE(n_rays > 0 && n_images > 0, "n_rays and n_images must be positive");
	cudaStream_t s = (cudaStream_t)stream;
	if (cudaMemsetAsync(d_scan_ws, 0, ngp_scan_workspace_bytes(n_rays), s) != cudaSuccess) { set_error("ngp_generate_training_samples: memset failed"); return -2; }
	Pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
	ScanState scan{(unsigned long long*)d_scan_ws};
	NGP_LAUNCH("generate_training_samples_kernel", s, generate_training_samples_kernel<<<div_up(n_rays, 128), 128, 0, s>>>(*cfg, *cam, n_rays, n_rays_shift, n_coords, n_images, d_xforms,
		(const uint2*)d_images, d_bitfield, snap_to_pixel_centers, rng, d_rays, (float4*)d_rays_rgba, d_numsteps, (Coordinate*)d_coords,
		d_counters, d_per_ray_n, scan));
	return check_launch("generate_training_samples");
}
