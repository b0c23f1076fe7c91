// Occupancy-grid ray marching primitives shared by the training and inference sampler kernels.
// Behaviour follows internal/sampler/occupancy_sampler.cu:207-308 and bounding_box.h:151-197; the operation order is the
// one documented in DESIGN.md (explicit __fmaf_rn, nothing else contracted; file compiled with -fmad=false).
#pragma once
#include "common.cuh"
#include "lattice_hop.h"

namespace ngp {

struct V3 { float x, y, z; };

__device__ __forceinline__ float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
__device__ __forceinline__ float calc_dt(float t, const ngp_sampler_cfg& c) { return clampf(t * c.cone_angle, c.min_step, c.max_step); }

__device__ __forceinline__ bool aabb_contains(const ngp_sampler_cfg& c, V3 p) {
	return p.x >= c.aabb_min && p.x <= c.aabb_max && p.y >= c.aabb_min && p.y <= c.aabb_max && p.z >= c.aabb_min && p.z <= c.aabb_max;
}

__device__ __forceinline__ void swapf(float& a, float& b) { float t = a; a = b; b = t; }

// slab test, returns tmin; miss -> 1e5 (bounding_box.h:151-186)
__device__ __forceinline__ float ray_intersect_tmin(const ngp_sampler_cfg& c, V3 o, V3 d) {
	float tmin = (c.aabb_min - o.x) / d.x, tmax = (c.aabb_max - o.x) / d.x;
	if (tmin > tmax) swapf(tmin, tmax);
	float tymin = (c.aabb_min - o.y) / d.y, tymax = (c.aabb_max - o.y) / d.y;
	if (tymin > tymax) swapf(tymin, tymax);
	if (tmin > tymax || tymin > tmax) return 100000.0f;
	if (tymin > tmin) tmin = tymin;
	if (tymax < tmax) tmax = tymax;
	float tzmin = (c.aabb_min - o.z) / d.z, tzmax = (c.aabb_max - o.z) / d.z;
	if (tzmin > tzmax) swapf(tzmin, tzmax);
	if (tmin > tzmax || tzmin > tmax) return 100000.0f;
	if (tzmin > tmin) tmin = tzmin;
	return tmin;
}

__device__ __forceinline__ int frexp_exponent(float v) { int e; frexpf(v, &e); return e; }

__device__ __forceinline__ int mip_from_pos(V3 p, uint32_t n_cascades) {
	const float m = fmaxf(fmaxf(fabsf(p.x - 0.5f), fabsf(p.y - 0.5f)), fabsf(p.z - 0.5f));
	return min((int)n_cascades - 1, max(0, frexp_exponent(m) + 1));
}

__device__ __forceinline__ int mip_from_dt(float dt, V3 p, const ngp_sampler_cfg& c) {
	const int mip = mip_from_pos(p, c.n_cascades);
	dt *= (float)(2 * c.grid_size);
	if (dt < 1.f) return mip;
	return min((int)c.n_cascades - 1, max(frexp_exponent(dt), mip));
}

__device__ __forceinline__ uint32_t cascaded_grid_idx_at(V3 p, uint32_t mip, uint32_t G) {
	const float s = scalbnf(1.0f, -(int)mip);
	const float g = (float)G;
	const int ix = min(max((int)(((p.x - 0.5f) * s + 0.5f) * g), 0), (int)G - 1);
	const int iy = min(max((int)(((p.y - 0.5f) * s + 0.5f) * g), 0), (int)G - 1);
	const int iz = min(max((int)(((p.z - 0.5f) * s + 0.5f) * g), 0), (int)G - 1);
	return morton3D(ix, iy, iz);
}

__device__ __forceinline__ bool grid_occupied(V3 p, const uint8_t* __restrict__ bits, uint32_t mip, uint32_t G) {
	const uint32_t idx = cascaded_grid_idx_at(p, mip, G);
	return __ldg(bits + idx / 8 + (size_t)(G * G * G) * mip / 8) & (1u << (idx % 8));
}

__device__ __forceinline__ float distance_to_next_voxel(V3 pos, V3 dir, V3 idir, uint32_t res) {
	const float r = (float)res;
	const float px = r * pos.x, py = r * pos.y, pz = r * pos.z;
	const float tx = (floorf(px + 0.5f + 0.5f * copysignf(1.f, dir.x)) - px) * idir.x;
	const float ty = (floorf(py + 0.5f + 0.5f * copysignf(1.f, dir.y)) - py) * idir.y;
	const float tz = (floorf(pz + 0.5f + 0.5f * copysignf(1.f, dir.z)) - pz) * idir.z;
	const float t = fminf(fminf(tx, ty), tz);
	return fmaxf(t / r, 0.0f);
}

__device__ __forceinline__ float advance_to_next_voxel(float t, const ngp_sampler_cfg& c, V3 pos, V3 dir, V3 idir, uint32_t res) {
	const float t_target = t + distance_to_next_voxel(pos, dir, idir, res);
	do { t += calc_dt(t, c); } while (t < t_target);
	return t;
}

// The same hop with the closed form of lattice_hop.h when the step is constant (cone_angle 0) and, for a power-of-two grid, the two
// divisions by the resolution written as exact multiplications.  Bit-identical to advance_to_next_voxel by construction: the
// closed form either reproduces the loop or declines.
__device__ __forceinline__ float advance_to_next_voxel_fast(float t, const ngp_sampler_cfg& c, bool const_dt, V3 pos, V3 dir, V3 idir, uint32_t res) {
	float dist;
	if ((res & (res - 1u)) == 0u) {
		const float r = (float)res, ir = 1.0f / r;            // exact: r is a power of two
		const float px = r * pos.x, py = r * pos.y, pz = r * pos.z;
		const float tx = (floorf(px + 0.5f + 0.5f * copysignf(1.f, dir.x)) - px) * idir.x;
		const float ty = (floorf(py + 0.5f + 0.5f * copysignf(1.f, dir.y)) - py) * idir.y;
		const float tz = (floorf(pz + 0.5f + 0.5f * copysignf(1.f, dir.z)) - pz) * idir.z;
		dist = fmaxf(fminf(fminf(tx, ty), tz) * ir, 0.0f);
	} else {
		dist = distance_to_next_voxel(pos, dir, idir, res);
	}
	const float t_target = t + dist;
	float hopped;
	if (const_dt && ngp_lattice_hop(t, t_target, c.min_step, &hopped)) return hopped;
	do { t += calc_dt(t, c); } while (t < t_target);
	return t;
}

__device__ __forceinline__ V3 ray_at(V3 o, V3 d, float t) { return {__fmaf_rn(t, d.x, o.x), __fmaf_rn(t, d.y, o.y), __fmaf_rn(t, d.z, o.z)}; }

__device__ __forceinline__ float warp_dt(float dt, const ngp_sampler_cfg& c) {
	const float max_stepsize = c.min_step * (float)(1u << (c.n_cascades - 1));
	return (dt - c.min_step) / (max_stepsize - c.min_step);
}

__device__ __forceinline__ Coordinate make_coord(V3 pos, V3 dir, float dt, const ngp_sampler_cfg& c) {
	const float diag = c.aabb_max - c.aabb_min;
	Coordinate o;
	o.px = (pos.x - c.aabb_min) / diag; o.py = (pos.y - c.aabb_min) / diag; o.pz = (pos.z - c.aabb_min) / diag;
	o.dt = warp_dt(dt, c);
	o.dx = (dir.x + 1.f) * 0.5f; o.dy = (dir.y + 1.f) * 0.5f; o.dz = (dir.z + 1.f) * 0.5f;
	return o;
}

// xf: 3x4 column-major camera-to-world; pixel_to_ray (internal/utils/common_device.h:130-152)
__device__ __forceinline__ void make_ray(const float* __restrict__ xf, float u, float v, const ngp_camera_cfg& cam, V3& o, V3& d) {
	const float cx = (u - cam.ppx) * (float)cam.width / cam.fx, cy = (v - cam.ppy) * (float)cam.height / cam.fy, cz = 1.0f;
	o = {xf[9], xf[10], xf[11]};
	// Eigen's fixed-size reductions associate as p0 + (p1 + p2) and nvcc contracts both sums (cz == 1: p2 is exact);
	// verified bit-for-bit against rays dumped from the reference's own sm_100 build
	float dx = __fmaf_rn(xf[0], cx, __fmaf_rn(xf[3], cy, xf[6] * cz));
	float dy = __fmaf_rn(xf[1], cx, __fmaf_rn(xf[4], cy, xf[7] * cz));
	float dz = __fmaf_rn(xf[2], cx, __fmaf_rn(xf[5], cy, xf[8] * cz));
	const float n2 = __fmaf_rn(dx, dx, __fmaf_rn(dy, dy, dz * dz));
	if (n2 > 0.f) { const float n = sqrtf(n2); dx /= n; dy /= n; dz /= n; }
	d = {dx, dy, dz};
}

} // namespace ngp
