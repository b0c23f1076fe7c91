// TEST INFRASTRUCTURE ONLY (see oracle_common.h).  Pinned to golden vectors from the reference's CUDA build (tests/golden/).
// CPU restatement of the occupancy sampler (internal/sampler/occupancy_sampler.cu):
//   mark_untrained_density_grid :18-61     generate_grid_samples_nerf_nonuniform :63-105
//   splat / ema                 :107-158   grid_to_bitfield / bitfield_max_pool  :160-205
//   calc_dt, mip_from_*, cascaded_grid_idx_at, distance/advance_to_next_voxel   :207-308
//   generate_training_samples_nerf :310-449   init_rays / advance_pos :451-550
//   generate_coords_from_rays_regularly_kernel :552-600
//   BoundingBox::ray_intersect / contains : internal/sampler/bounding_box.h:151-197
// Operation order is explicit (fmaf where the CUDA kernels use __fmaf_rn; nothing else contracts).
// Atomic-ordered compaction of the reference is restated in ray order (deterministic).
#include "oracle_common.h"
#include <omp.h>

using namespace orc;

namespace {

struct SamplerCfg {
	float aabb_min, aabb_max;          // cube [min,max]^3
	uint32_t grid_size, n_cascades;
	float cone_angle, min_step, max_step, near_distance;
	uint32_t max_steps;
};

static inline float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }
static inline float calc_dt(float t, const SamplerCfg& c) { return clampf(t * c.cone_angle, c.min_step, c.max_step); }

static inline bool contains(const SamplerCfg& c, V3 p) {
	return p.x >= c.aabb_min && p.x <= c.aabb_max && p.y >= c.aabb_min && p.y <= c.aabb_max && p.z >= c.aabb_min && p.z <= c.aabb_max;
}

static inline void swapf(float& a, float& b) { float t = a; a = b; b = t; }

// returns tmin (tmax unused by callers); miss -> 1e5
static inline float ray_intersect_tmin(const SamplerCfg& c, V3 o, V3 d) {
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

static inline int mip_from_pos(V3 p, uint32_t n_cascades) {
	int e;
	float m = fmaxf(fmaxf(fabsf(p.x - 0.5f), fabsf(p.y - 0.5f)), fabsf(p.z - 0.5f));
	frexpf(m, &e);
	return std::min((int)n_cascades - 1, std::max(0, e + 1));
}

static inline int mip_from_dt(float dt, V3 p, const SamplerCfg& c) {
	int mip = mip_from_pos(p, c.n_cascades);
	dt *= 2 * c.grid_size;
	if (dt < 1.f) return mip;
	int e;
	frexpf(dt, &e);
	return std::min((int)c.n_cascades - 1, std::max(e, mip));
}

static inline uint32_t cascaded_grid_idx_at(V3 p, uint32_t mip, uint32_t G) {
	float s = scalbnf(1.0f, -(int)mip);
	float q[3] = {(p.x - 0.5f) * s + 0.5f, (p.y - 0.5f) * s + 0.5f, (p.z - 0.5f) * s + 0.5f};
	int i[3];
	for (int d = 0; d < 3; ++d) i[d] = std::min(std::max((int)(q[d] * (float)G), 0), (int)G - 1);
	return morton3D(i[0], i[1], i[2]);
}

static inline bool occupied(V3 p, const uint8_t* bits, uint32_t mip, uint32_t G) {
	uint32_t idx = cascaded_grid_idx_at(p, mip, G);
	return bits[idx / 8 + (size_t)(G * G * G) * mip / 8] & (1u << (idx % 8));
}

static inline float distance_to_next_voxel(V3 pos, V3 dir, V3 idir, uint32_t res) {
	float r = (float)res;
	float px = r * pos.x, py = r * pos.y, pz = r * pos.z;
	float tx = (floorf(px + 0.5f + 0.5f * copysignf(1.f, dir.x)) - px) * idir.x;
	float ty = (floorf(py + 0.5f + 0.5f * copysignf(1.f, dir.y)) - py) * idir.y;
	float tz = (floorf(pz + 0.5f + 0.5f * copysignf(1.f, dir.z)) - pz) * idir.z;
	float t = fminf(fminf(tx, ty), tz);
	return fmaxf(t / r, 0.0f);
}

static inline float advance_to_next_voxel(float t, const SamplerCfg& c, V3 pos, V3 dir, V3 idir, uint32_t res) {
	float t_target = t + distance_to_next_voxel(pos, dir, idir, res);
	do { t += calc_dt(t, c); } while (t < t_target);
	return t;
}

static inline V3 at(V3 o, V3 d, float t) { return {fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z)}; }

static inline float warp_dt(float dt, const SamplerCfg& c) {
	float max_stepsize = c.min_step * (float)(1u << (c.n_cascades - 1));
	return (dt - c.min_step) / (max_stepsize - c.min_step);
}

static inline Coordinate make_coord(V3 pos, V3 dir, float dt, const SamplerCfg& c) {
	float diag = c.aabb_max - c.aabb_min;
	Coordinate o;
	o.px = (pos.x - c.aabb_min) / diag; o.py = (pos.y - c.aabb_min) / diag; o.pz = (pos.z - c.aabb_min) / diag;
	o.dt = warp_dt(dt, c);
	o.dx = (dir.x + 1.f) * 0.5f; o.dy = (dir.y + 1.f) * 0.5f; o.dz = (dir.z + 1.f) * 0.5f;
	return o;
}

// xform: 12 floats, column-major 3x4 (cols 0..2 rotation, col 3 origin)
static inline void make_ray(const float* xf, float u, float v, float ppx, float ppy, int W, int H, float fx, float fy, V3& o, V3& d) {
	float cx = (u - ppx) * (float)W / fx, cy = (v - ppy) * (float)H / fy, cz = 1.0f;
	o = {xf[9], xf[10], xf[11]};
	// Eigen's fixed-size reductions associate as p0 + (p1 + p2) and nvcc contracts both sums; cz == 1 makes p2 exact.
	// Confirmed bit-for-bit against rays dumped from the reference's sm_100 build (tests/test_gpu_reference_cuda.py).
	float dx = fmaf(xf[0], cx, fmaf(xf[3], cy, xf[6] * cz));
	float dy = fmaf(xf[1], cx, fmaf(xf[4], cy, xf[7] * cz));
	float dz = fmaf(xf[2], cx, fmaf(xf[5], cy, xf[8] * cz));
	float n2 = fmaf(dx, dx, fmaf(dy, dy, dz * dz));
	if (n2 > 0.f) { float n = sqrtf(n2); dx /= n; dy /= n; dz /= n; }
	d = {dx, dy, dz};
}

static inline SamplerCfg make_cfg(const float* f, const uint32_t* u) {
	SamplerCfg c;
	c.aabb_min = f[0]; c.aabb_max = f[1]; c.cone_angle = f[2]; c.min_step = f[3]; c.max_step = f[4]; c.near_distance = f[5];
	c.grid_size = u[0]; c.n_cascades = u[1]; c.max_steps = u[2];
	return c;
}

} // namespace

#include "sampler_grid.inl"
#include "sampler_march.inl"
