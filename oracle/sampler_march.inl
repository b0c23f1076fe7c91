// TEST INFRASTRUCTURE ONLY (see oracle_common.h).  Included by sampler.cpp.
// Ray generation + occupancy march: occupancy_sampler.cu:310-600; pixel_to_ray common_device.h:130-152;
// low-discrepancy jitter random_val.cuh:142-278.

namespace {

// Sobol direction numbers for dimensions 0 and 1 (the only ones the path uses): v0[b] = 2^(31-b),
// v1[b] = v1[b-1] ^ (v1[b-1] >> 1)
static inline uint32_t sobol(uint32_t index, uint32_t dim) {
	uint32_t X = 0, v = 0x80000000u;
	for (uint32_t bit = 0; bit < 32; ++bit) {
		uint32_t dirn = dim == 0 ? (0x80000000u >> bit) : v;
		if ((index >> bit) & 1) X ^= dirn;
		v = v ^ (v >> 1);
	}
	return X;
}
static inline uint32_t hash_combine(uint32_t seed, uint32_t v) { return seed ^ (v + (seed << 6) + (seed >> 2)); }
static inline uint32_t reverse_bits(uint32_t x) {
	x = ((x & 0xaaaaaaaau) >> 1) | ((x & 0x55555555u) << 1);
	x = ((x & 0xccccccccu) >> 2) | ((x & 0x33333333u) << 2);
	x = ((x & 0xf0f0f0f0u) >> 4) | ((x & 0x0f0f0f0fu) << 4);
	x = ((x & 0xff00ff00u) >> 8) | ((x & 0x00ff00ffu) << 8);
	return (x >> 16) | (x << 16);
}
static inline uint32_t laine_karras(uint32_t x, uint32_t seed) {
	x += seed; x ^= x * 0x6c50b47cu; x ^= x * 0xb82f1e52u; x ^= x * 0xc7afe638u; x ^= x * 0x8d22f6e6u;
	return x;
}
static inline uint32_t nus_base2(uint32_t x, uint32_t seed) { return reverse_bits(laine_karras(reverse_bits(x), seed)); }
static inline float ld_random_val(uint32_t index, uint32_t seed, uint32_t dim) {
	const float S = float(1.0 / (1ull << 32));
	index = nus_base2(index, seed);
	return (float)nus_base2(sobol(index, dim), hash_combine(seed, dim)) * S;
}
static inline void ld_random_val_2d(uint32_t index, uint32_t seed, float& a, float& b) {
	const float S = float(1.0 / (1ull << 32));
	index = nus_base2(index, seed);
	a = (float)nus_base2(sobol(index, 0), hash_combine(seed, 0)) * S;
	b = (float)nus_base2(sobol(index, 1), hash_combine(seed, 1)) * S;
}
static inline float fractf(float x) { return x - floorf(x); }

} // namespace

extern "C" {

float orc_ld_random_val(uint32_t index, uint32_t seed, uint32_t dim) { return ld_random_val(index, seed, dim); }

// Training rays + samples (:310-449), compaction in ray order.
// cfg_f: aabb_min,aabb_max,cone_angle,min_step,max_step,near ; cfg_u: grid_size,n_cascades,max_steps
// images: fp16 RGBA bits [n_images][H][W][4].  Outputs: rays [R][6], rgba [R][4], numsteps [R][2] (count, base),
// coords [n_coords][7]; counters[0] = rays kept, counters[1] = sum of per-ray counts (including dropped rays).
// per_ray_n (optional, [n_rays]): count of every generated ray (0 = no sample), for bit-exact count checks.
void orc_generate_training_samples(uint32_t n_rays, uint32_t n_rays_shift, uint32_t n_coords, uint32_t n_images, int W, int H,
                                   float fx, float fy, float ppx, float ppy, const float* xforms, const uint16_t* images,
                                   const float* cfg_f, const uint32_t* cfg_u, const uint8_t* bits, int snap,
                                   uint64_t rng_state, uint64_t rng_inc,
                                   float* rays_out, float* rgba_out, uint32_t* numsteps_out, float* coords_out,
                                   uint32_t* counters, uint32_t* per_ray_n) {
	const SamplerCfg c = make_cfg(cfg_f, cfg_u);
	struct Tmp { V3 o, d; float startt, u, v; uint32_t n, img; };
	std::vector<Tmp> tmp(n_rays);
	#pragma omp parallel for schedule(dynamic, 64)
	for (int64_t ii = 0; ii < (int64_t)n_rays; ++ii) {
		uint32_t i = (uint32_t)ii;
		Tmp& r = tmp[i];
		r.img = (((n_rays_shift + i) * n_images) / n_rays) % n_images;
		Pcg32 rng; rng.state = rng_state; rng.inc = rng_inc;
		rng.advance((int64_t)((uint64_t)i * 8));
		float u = rng.next_float(), v = rng.next_float();
		if (snap) {
			u = ((float)std::min(std::max((int)(u * (float)W), 0), W - 1) + 0.5f) / (float)W;
			v = ((float)std::min(std::max((int)(v * (float)H), 0), H - 1) + 0.5f) / (float)H;
		}
		r.u = u; r.v = v;
		make_ray(xforms + (size_t)r.img * 12, u, v, ppx, ppy, W, H, fx, fy, r.o, r.d);
		float tmin = fmaxf(ray_intersect_tmin(c, r.o, r.d), c.near_distance);
		float startt = fmaf(calc_dt(tmin, c), rng.next_float(), tmin);
		r.startt = startt;
		V3 idir = {1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z};
		uint32_t j = 0; float t = startt; V3 pos;
		while (contains(c, pos = at(r.o, r.d, t)) && j < c.max_steps) {
			float dt = calc_dt(t, c);
			uint32_t mip = mip_from_dt(dt, pos, c);
			if (occupied(pos, bits, mip, c.grid_size)) { ++j; t += dt; }
			else t = advance_to_next_voxel(t, c, pos, r.d, idir, c.grid_size >> mip);
		}
		r.n = j;
		if (per_ray_n) per_ray_n[i] = j;
	}
	uint32_t total = 0, kept = 0;
	for (uint32_t i = 0; i < n_rays; ++i) {
		Tmp& r = tmp[i];
		if (r.n == 0) continue;
		uint32_t base = total; total += r.n;
		if (base + r.n > n_coords) continue;
		uint32_t ray_idx = kept++;
		float* ro = rays_out + (size_t)ray_idx * 6;
		ro[0] = r.o.x; ro[1] = r.o.y; ro[2] = r.o.z; ro[3] = r.d.x; ro[4] = r.d.y; ro[5] = r.d.z;
		int px = std::max(std::min((int)(r.u * (float)W), W - 1), 0), py = std::max(std::min((int)(r.v * (float)H), H - 1), 0);
		const uint16_t* pix = images + ((size_t)px + (size_t)py * W + (size_t)r.img * W * H) * 4;
		for (int k = 0; k < 4; ++k) rgba_out[(size_t)ray_idx * 4 + k] = h16from(pix[k]);
		numsteps_out[(size_t)ray_idx * 2 + 0] = r.n; numsteps_out[(size_t)ray_idx * 2 + 1] = base;
		V3 idir = {1.0f / r.d.x, 1.0f / r.d.y, 1.0f / r.d.z};
		Coordinate* out = (Coordinate*)coords_out + base;
		uint32_t j = 0; float t = r.startt; V3 pos;
		while (contains(c, pos = at(r.o, r.d, t)) && j < r.n) {
			float dt = calc_dt(t, c);
			uint32_t mip = mip_from_dt(dt, pos, c);
			if (occupied(pos, bits, mip, c.grid_size)) { out[j] = make_coord(pos, r.d, dt, c); ++j; t += dt; }
			else t = advance_to_next_voxel(t, c, pos, r.d, idir, c.grid_size >> mip);
		}
	}
	counters[0] = kept; counters[1] = total;
}

// Inference ray init + advance to first occupied cell (:451-550).  payloads: [W*H] RayPayload (40 B).
void orc_init_rays(int W, int H, float fx, float fy, float ppx, float ppy, const float* cam, uint32_t spp, int snap,
                   const float* cfg_f, const uint32_t* cfg_u, const uint8_t* bits, void* payloads_) {
	const SamplerCfg c = make_cfg(cfg_f, cfg_u);
	RayPayload* payloads = (RayPayload*)payloads_;
	float ox0, oy0, ox1, oy1;
	ld_random_val_2d(0, 0xdeadbeefu, ox0, oy0);
	ld_random_val_2d(snap ? 0 : spp, 0xdeadbeefu, ox1, oy1);
	const float offx = fractf(0.5f - ox0 + ox1), offy = fractf(0.5f - oy0 + oy1);
	#pragma omp parallel for schedule(dynamic, 256)
	for (int64_t ii = 0; ii < (int64_t)W * H; ++ii) {
		uint32_t idx = (uint32_t)ii, x = idx % W, y = idx / W;
		RayPayload& p = payloads[idx];
		V3 o, d;
		make_ray(cam, ((float)x + offx) / (float)W, ((float)y + offy) / (float)H, ppx, ppy, W, H, fx, fy, o, d);
		float t = fmaxf(ray_intersect_tmin(c, o, d), 0.05f) + 1e-6f;
		std::memset(&p, 0, sizeof(p));
		p.ox = o.x; p.oy = o.y; p.oz = o.z;
		if (!contains(c, at(o, d, t))) { p.alive = 0; continue; }
		p.dx = d.x; p.dy = d.y; p.dz = d.z; p.t = t; p.idx = idx; p.n_steps = 0; p.alive = 1;
		// advance_pos_nerf
		V3 idir = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
		float dt = calc_dt(t, c);
		t = fmaf(ld_random_val(spp, idx * 786433u, 0), dt, t);
		V3 pos;
		while (true) {
			if (!contains(c, pos = at(o, d, t))) { p.alive = 0; break; }
			dt = calc_dt(t, c);
			uint32_t mip = mip_from_dt(dt, pos, c);
			if (occupied(pos, bits, mip, c.grid_size)) break;
			t = advance_to_next_voxel(t, c, pos, d, idir, c.grid_size >> mip);
		}
		p.t = t;
	}
}

// (:552-600) up to n_steps occupied samples per alive ray at coords[i*n_steps + j]
void orc_generate_coords_from_rays(uint32_t n_rays, uint32_t n_steps, const float* cfg_f, const uint32_t* cfg_u,
                                   const uint8_t* bits, void* payloads_, float* coords_out) {
	const SamplerCfg c = make_cfg(cfg_f, cfg_u);
	RayPayload* payloads = (RayPayload*)payloads_;
	#pragma omp parallel for schedule(dynamic, 256)
	for (int64_t ii = 0; ii < (int64_t)n_rays; ++ii) {
		RayPayload& p = payloads[ii];
		if (!p.alive) continue;
		V3 o = {p.ox, p.oy, p.oz}, d = {p.dx, p.dy, p.dz}, idir = {1.0f / d.x, 1.0f / d.y, 1.0f / d.z};
		float t = p.t;
		bool done = false;
		for (uint32_t j = 0; j < n_steps && !done; ++j) {
			V3 pos; float dt = 0.f;
			while (true) {
				if (!contains(c, pos = at(o, d, t))) { p.n_steps = j; done = true; break; }
				dt = calc_dt(t, c);
				uint32_t mip = mip_from_dt(dt, pos, c);
				if (occupied(pos, bits, mip, c.grid_size)) break;
				t = advance_to_next_voxel(t, c, pos, d, idir, c.grid_size >> mip);
			}
			if (done) break;
			((Coordinate*)coords_out)[(size_t)ii * n_steps + j] = make_coord(pos, d, dt, c);
			t += dt;
		}
		if (!done) { p.t = t; p.n_steps = n_steps; }
	}
}

} // extern "C"
