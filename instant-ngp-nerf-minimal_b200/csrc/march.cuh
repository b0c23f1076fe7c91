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

// ---- the per-point pieces of the scalar march (one visited lattice point per call), shared by the inference kernels
// (march_infer.cu) and the training march (march_train.cu) ----------------------------------------------------------------

// mip_from_dt (occupancy_sampler.cu:225-240) with frexp's exponent read from the float's bits: frexpf(m) = (f, e) with
// e = biased exponent - 126 for normal m and e = 0 for m == 0; subnormal m gives e <= -126 either way, clamped to cascade 0
__device__ __forceinline__ int exponent_of(float m) { return m == 0.f ? 0 : (int)(__float_as_uint(m) >> 23 & 0xffu) - 126; }
__device__ __forceinline__ uint32_t mip_at(float dt, V3 p, const ngp_sampler_cfg& c) {
	const float m = fmaxf(fmaxf(fabsf(p.x - 0.5f), fabsf(p.y - 0.5f)), fabsf(p.z - 0.5f));
	const int mip = min((int)c.n_cascades - 1, max(0, exponent_of(m) + 1));
	dt *= (float)(2 * c.grid_size);
	if (dt < 1.f) return (uint32_t)mip;
	return (uint32_t)min((int)c.n_cascades - 1, max(exponent_of(dt), mip));
}

// the cell around p in cascade mip: its coordinates in cells (float, before truncation), its Morton index and its occupancy bit.
// The Morton interleave comes from a 256-entry shared table when the grid is small enough (every grid the configurations use).
// `known` (optional): packed coordinates of a cell already known to be OCCUPIED (the previous sample's); a probe that lands in it
// again -- about four of five samples do -- skips the table look-ups and the bitfield load.  0xffffffff = none.
struct CellProbe { float gx, gy, gz; uint32_t idx; bool occupied; };
__device__ __forceinline__ CellProbe probe_cell(V3 p, uint32_t mip, uint32_t G, const uint32_t* __restrict__ spread, const uint8_t* __restrict__ bits, uint32_t* known = nullptr) {
	const float s = __uint_as_float((127u - mip) << 23), g = (float)G;          // scalbnf(1, -mip)
	CellProbe cp;
	cp.gx = ((p.x - 0.5f) * s + 0.5f) * g; cp.gy = ((p.y - 0.5f) * s + 0.5f) * g; cp.gz = ((p.z - 0.5f) * s + 0.5f) * g;
	const int ix = min(max((int)cp.gx, 0), (int)G - 1), iy = min(max((int)cp.gy, 0), (int)G - 1), iz = min(max((int)cp.gz, 0), (int)G - 1);
	const uint32_t key = known && G <= 512 ? ((uint32_t)ix | ((uint32_t)iy << 9) | ((uint32_t)iz << 18) | (mip << 27)) : 0xffffffffu;
	if (known && key == *known && key != 0xffffffffu) { cp.idx = 0; cp.occupied = true; return cp; }
	cp.idx = spread ? (spread[ix] | (spread[iy] << 1) | (spread[iz] << 2)) : morton3D(ix, iy, iz);
	cp.occupied = __ldg(bits + cp.idx / 8 + (size_t)(G * G * G) * mip / 8) & (1u << (cp.idx % 8));
	if (known) *known = cp.occupied ? key : 0xffffffffu;
	return cp;
}

// Crossing a whole empty 8x8x8 block of cells in one hop, with the reference's result.
//
// The reference visits (about) one lattice point per empty cell: from a visited point v it goes to the first lattice point at or
// behind  tau = fl(v.t + exit distance of v's cell).  tau differs from the true time e at which the ray crosses that cell face by a
// rounding error |tau - e| <= delta, so whenever no lattice point lies within delta of e the landing point is L(e), the first
// lattice point behind e -- whatever v was.  Now let e* be the time the ray leaves the block B around the current point, all of B
// empty.  The reference's last visited point before e* sits in some (empty) cell of B; if its own hop aims at a face before e* it
// nevertheless lands behind e* (it is the LAST point before e*), so no lattice point lies between its target and e*, and its landing
// point is L(e*) again.  Hence, if no lattice point lies within delta + delta' of our own estimate of e* (delta' its error), hopping
// straight to L(e*) reproduces a point the reference visits, and every point it visited in between lay in empty cells: the march
// continues from identical state.  Otherwise the hop is declined and the cell-by-cell march goes on (it re-tries in the next cell).
//   delta, delta' <= 2 ulp(box coordinate) |1 / d_axis| + ulp(t) each (position fma, the products with 1/d, the sum with t); the
//   margin used is 8 (ulp(coordinate) |1 / d_axis| + ulp(t)) plus 1e-3 of a step.
// Needs the constant-step lattice (L.inv_dtq != 0) and stays inside its binade; blocks never straddle a cascade shell (shell faces
// are at multiples of 8 cells for grid_size >= 32).
__device__ __forceinline__ bool coarse_hop(const ngp_sampler_cfg& c, const ngp_lattice& L, float& t, const CellProbe& cp, uint32_t mip, V3 d, V3 idir, float coord_ulp) {
	if (!c.coarse_bitfield || L.inv_dtq == 0.f) return false;
	const uint32_t G = c.grid_size, block = cp.idx >> 9;
	if (__ldg(c.coarse_bitfield + (size_t)(G * G * G / 4096) * mip + (block >> 3)) & (1u << (block & 7u))) return false;
	if (!(cp.gx >= 0.f && cp.gx < (float)G && cp.gy >= 0.f && cp.gy < (float)G && cp.gz >= 0.f && cp.gz < (float)G)) return false;
	// exit of the block, in cells along each axis, as a time
	const float bx = floorf(cp.gx * 0.125f) * 8.f, by = floorf(cp.gy * 0.125f) * 8.f, bz = floorf(cp.gz * 0.125f) * 8.f;
	const float tx = d.x > 0.f ? (bx + 8.f - cp.gx) * idir.x : d.x < 0.f ? (bx - cp.gx) * idir.x : INFINITY;
	const float ty = d.y > 0.f ? (by + 8.f - cp.gy) * idir.y : d.y < 0.f ? (by - cp.gy) * idir.y : INFINITY;
	const float tz = d.z > 0.f ? (bz + 8.f - cp.gz) * idir.z : d.z < 0.f ? (bz - cp.gz) * idir.z : INFINITY;
	float tmin = tx, ia = fabsf(idir.x);
	if (ty < tmin) { tmin = ty; ia = fabsf(idir.y); }
	if (tz < tmin) { tmin = tz; ia = fabsf(idir.z); }
	const float dist = tmin * (__uint_as_float((127u + mip) << 23) / (float)G);      // cells -> world units: 2^mip / G (exact: G is a power of two)
	if (!(dist > 0.f) || !(dist < 1.0e30f)) return false;
	const float x = dist * L.inv_dtq, fl = floorf(x), fr = x - fl;
	const float margin = 8.f * (coord_ulp * ia + L.top * 5.9604645e-8f) * L.inv_dtq + 1.0e-3f;
	if (!(fr > margin && fr < 1.f - margin) || !(fl < 8388607.f)) return false;
	const float tc = fmaf(fl + 1.f, L.dtq, t);
	if (!(tc < L.top)) return false;
	t = tc;
	return true;
}

// advance_to_next_voxel (occupancy_sampler.cu:284-294): distance to the cell's exit, then the first lattice point at or behind it --
// in closed form where lattice_hop.h can prove it, else with the reference's loop
__device__ __forceinline__ void cell_hop(const ngp_sampler_cfg& c, ngp_lattice& L, float& t, V3 pos, V3 idir, V3 half_sign, uint32_t mip, bool const_dt) {
	const uint32_t res = c.grid_size >> mip;
	const float rr = (float)res;
	const float qx = rr * pos.x, qy = rr * pos.y, qz = rr * pos.z;
	const float tx = (floorf(qx + 0.5f + half_sign.x) - qx) * idir.x;
	const float ty = (floorf(qy + 0.5f + half_sign.y) - qy) * idir.y;
	const float tz = (floorf(qz + 0.5f + half_sign.z) - qz) * idir.z;
	const float tmin = fminf(fminf(tx, ty), tz);
	// t / res as a multiplication when res is a power of two (exact)
	const float dist = fmaxf((res & (res - 1u)) == 0u ? tmin * __uint_as_float((127u - (31u - (uint32_t)__clz((int)res))) << 23) : tmin / rr, 0.0f);
	const float t_target = t + dist;
	float hopped;
	if (ngp_lattice_hop_cached(&L, t, t_target, &hopped)) { t = hopped; return; }
	do { t += calc_dt(t, c); } while (t < t_target);
	ngp_lattice_enter(&L, t, c.min_step, const_dt);
}

} // namespace ngp
