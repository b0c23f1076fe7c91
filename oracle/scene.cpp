// TEST INFRASTRUCTURE ONLY (see oracle_common.h).  CPU twin of the procedural benchmark scene
// (instant-ngp-nerf-minimal_b200/csrc/synthetic_scene.cu) so the CPU arm of bench.py trains on the same kind of data
// without touching the GPU.  Fixture, not reference arithmetic: images agree with the CUDA generator to 8-bit rounding.
#include "oracle_common.h"
#include <omp.h>

using namespace orc;

namespace {
constexpr int kBlobs = 8;
struct Blob { float c[3]; float inv_s2[3]; float amp; float col[3]; float lobe[3]; };
struct Scene { Blob blobs[kBlobs]; float shell_c[3]; float shell_r, shell_t, shell_amp; float shell_col[3]; };

static Scene make_scene(uint32_t seed) {
	Scene s;
	Pcg32 rng{(uint64_t)seed};
	for (int i = 0; i < kBlobs; ++i) {
		Blob& b = s.blobs[i];
		float n2 = 0.f;
		for (int k = 0; k < 3; ++k) {
			b.c[k] = 0.3f + 0.4f * rng.next_float();
			const float sd = 0.025f + 0.05f * rng.next_float();
			b.inv_s2[k] = 1.f / (sd * sd);
			b.col[k] = 0.15f + 0.85f * rng.next_float();
			b.lobe[k] = rng.next_float() * 2.f - 1.f; n2 += b.lobe[k] * b.lobe[k];
		}
		for (int k = 0; k < 3; ++k) b.lobe[k] /= sqrtf(n2 > 0.f ? n2 : 1.f);
		b.amp = 40.f + 80.f * rng.next_float();
	}
	s.shell_c[0] = 0.5f; s.shell_c[1] = 0.5f; s.shell_c[2] = 0.34f;
	s.shell_r = 0.11f; s.shell_t = 0.006f; s.shell_amp = 400.f;
	s.shell_col[0] = 0.9f; s.shell_col[1] = 0.8f; s.shell_col[2] = 0.3f;
	return s;
}

static inline void scene_eval(const Scene& s, float x, float y, float z, float dx, float dy, float dz, float& sigma, float* rgb) {
	sigma = 0.f; rgb[0] = rgb[1] = rgb[2] = 0.f;
	for (int i = 0; i < kBlobs; ++i) {
		const Blob& b = s.blobs[i];
		const float ex = x - b.c[0], ey = y - b.c[1], ez = z - b.c[2];
		const float q = ex * ex * b.inv_s2[0] + ey * ey * b.inv_s2[1] + ez * ez * b.inv_s2[2];
		if (q < 18.f) {
			const float d = b.amp * expf(-0.5f * q);
			const float v = 0.65f + 0.35f * (dx * b.lobe[0] + dy * b.lobe[1] + dz * b.lobe[2]);
			sigma += d;
			rgb[0] += d * b.col[0] * v; rgb[1] += d * b.col[1] * v; rgb[2] += d * b.col[2] * v;
		}
	}
	const float ex = x - s.shell_c[0], ey = y - s.shell_c[1], ez = z - s.shell_c[2];
	const float r = sqrtf(ex * ex + ey * ey + ez * ez);
	if (fabsf(r - s.shell_r) < s.shell_t) {
		const float band = (((int)floorf(ey * 40.f) + (int)floorf(atan2f(ez, ex) * 6.f)) & 1) ? 1.0f : 0.35f;
		sigma += s.shell_amp;
		for (int k = 0; k < 3; ++k) rgb[k] += s.shell_amp * s.shell_col[k] * band;
	}
	if (sigma > 0.f) { const float inv = 1.f / sigma; rgb[0] *= inv; rgb[1] *= inv; rgb[2] *= inv; }
	for (int k = 0; k < 3; ++k) rgb[k] = fminf(fmaxf(rgb[k], 0.f), 1.f);
}
} // namespace

extern "C" {

// xforms: [n_views] 3x4 column-major camera-to-world in the unit-cube frame; out: rgba8 [n_views][H][W][4]
void orc_render_procedural_views(uint32_t seed, uint32_t n_views, int W, int H, float fx, float fy, const float* xforms, uint8_t* out) {
	const Scene scene = make_scene(seed);
	const int64_t n_pix = (int64_t)W * H;
	#pragma omp parallel for schedule(dynamic, 256)
	for (int64_t gid = 0; gid < n_pix * n_views; ++gid) {
		const uint32_t view = (uint32_t)(gid / n_pix), pix = (uint32_t)(gid % n_pix);
		const int px = pix % W, py = pix / W;
		const float* xf = xforms + (size_t)view * 12;
		const float cx = (((float)px + 0.5f) / (float)W - 0.5f) * (float)W / fx, cy = (((float)py + 0.5f) / (float)H - 0.5f) * (float)H / fy;
		float d[3] = {xf[0] * cx + xf[3] * cy + xf[6], xf[1] * cx + xf[4] * cy + xf[7], xf[2] * cx + xf[5] * cy + xf[8]};
		const float inv = 1.f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
		for (float& v : d) v *= inv;
		const float o[3] = {xf[9], xf[10], xf[11]};
		float t0 = 0.f, t1 = 1e9f;
		for (int k = 0; k < 3; ++k) {
			const float a = (0.f - o[k]) / d[k], b = (1.f - o[k]) / d[k];
			t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
		}
		float C[3] = {0, 0, 0}, T = 1.f;
		if (t1 > t0) {
			const float dt = 1.73205080757f / 1024.f;
			for (float t = t0 + 0.5f * dt; t < t1 && T > 1e-4f; t += dt) {
				float sigma, rgb[3];
				scene_eval(scene, o[0] + t * d[0], o[1] + t * d[1], o[2] + t * d[2], d[0], d[1], d[2], sigma, rgb);
				if (sigma <= 0.f) continue;
				const float alpha = 1.f - expf(-sigma * dt), w = alpha * T;
				for (int k = 0; k < 3; ++k) C[k] += w * rgb[k];
				T *= 1.f - alpha;
			}
		}
		const float A = 1.f - T;
		uint8_t* q = out + (size_t)gid * 4;
		for (int k = 0; k < 3; ++k) {
			const float straight = A > 1e-6f ? fminf(C[k] / A, 1.f) : 0.f;
			q[k] = (uint8_t)(fminf(fmaxf(linear_to_srgb(straight), 0.f), 1.f) * 255.f + 0.5f);
		}
		q[3] = (uint8_t)(fminf(fmaxf(A, 0.f), 1.f) * 255.f + 0.5f);
	}
}

// from_rgba32 (internal/utils/common_device.h:160-175): u8 sRGB straight alpha -> linear premultiplied fp16 bits
void orc_from_rgba32(uint64_t n_pixels, const uint8_t* in, uint16_t* out) {
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)n_pixels; ++i) {
		const uint8_t* p = in + (size_t)i * 4;
		const float a = (float)p[3] * (1.0f / 255.0f);
		for (int k = 0; k < 3; ++k) out[(size_t)i * 4 + k] = h16bits(srgb_to_linear((float)p[k] * (1.0f / 255.0f)) * a);
		out[(size_t)i * 4 + 3] = h16bits(a);
	}
}

// analytic optical thickness of the scene per occupancy cell (Morton order, cascade-major): a trained grid's stand-in so
// the CPU arm marches a realistic occupancy pattern from its first step
void orc_scene_density_grid(uint32_t seed, uint32_t G, uint32_t n_cascades, float min_step, float* grid) {
	const Scene scene = make_scene(seed);
	const uint32_t cells = G * G * G;
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)cells * n_cascades; ++i) {
		const uint32_t level = (uint32_t)(i / cells), idx = (uint32_t)(i % cells);
		const uint32_t x = morton3D_invert(idx), y = morton3D_invert(idx >> 1), z = morton3D_invert(idx >> 2);
		const float s = scalbnf(1.0f, (int)level);
		float best = 0.f;
		for (int k = 0; k < 8; ++k) {   // the 8 sub-cell centres: a cell is as dense as its densest probe
			const float fx = ((float)x + 0.25f + 0.5f * (k & 1)) / (float)G, fy = ((float)y + 0.25f + 0.5f * ((k >> 1) & 1)) / (float)G, fz = ((float)z + 0.25f + 0.5f * (k >> 2)) / (float)G;
			float sigma, rgb[3];
			scene_eval(scene, (fx - 0.5f) * s + 0.5f, (fy - 0.5f) * s + 0.5f, (fz - 0.5f) * s + 0.5f, 0.f, 0.f, 1.f, sigma, rgb);
			best = fmaxf(best, sigma);
		}
		grid[i] = best * min_step;
	}
}

} // extern "C"
