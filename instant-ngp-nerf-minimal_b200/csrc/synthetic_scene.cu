// Procedural benchmark scene (SURVEY.md 8d "C2/C3 synthetic scene"): an analytic density + view-dependent colour field
// inside the unit cube, ray-marched into 8-bit sRGB RGBA views of nerf_synthetic shape.  There is no network on the GPU
// box, so this replaces the Blender renders; the images then go through the same from_rgba32 conversion as the
// reference's PNG loader (internal/utils/common_device.h:160-175).  Fixture code: runs once at load, never timed.
#include "common.cuh"

namespace ngp {

struct Blob { float c[3]; float inv_s2[3]; float amp; float col[3]; float lobe[3]; };
struct Scene { Blob blobs[NGP_SCENE_BLOBS]; float shell_c[3]; float shell_r, shell_t, shell_amp; float shell_col[3]; };

__device__ __forceinline__ void scene_eval(const Scene& s, float x, float y, float z, float dx, float dy, float dz, float& sigma, float* rgb) {
	sigma = 0.f; rgb[0] = rgb[1] = rgb[2] = 0.f;
	#pragma unroll
	for (int i = 0; i < NGP_SCENE_BLOBS; ++i) {
		const Blob& b = s.blobs[i];
		const float ex = x - b.c[0], ey = y - b.c[1], ez = z - b.c[2];
		const float q = ex * ex * b.inv_s2[0] + ey * ey * b.inv_s2[1] + ez * ez * b.inv_s2[2];
		if (q < 18.f) {
			const float d = b.amp * __expf(-0.5f * q);
			const float v = 0.65f + 0.35f * (dx * b.lobe[0] + dy * b.lobe[1] + dz * b.lobe[2]);
			sigma += d;
			rgb[0] += d * b.col[0] * v; rgb[1] += d * b.col[1] * v; rgb[2] += d * b.col[2] * v;
		}
	}
	{
		const float ex = x - s.shell_c[0], ey = y - s.shell_c[1], ez = z - s.shell_c[2];
		const float r = sqrtf(ex * ex + ey * ey + ez * ez);
		if (fabsf(r - s.shell_r) < s.shell_t) {
			// hard-edged shell with a banded albedo so fine hash levels have something to learn
			const float band = (((int)floorf(ey * 40.f) + (int)floorf(atan2f(ez, ex) * 6.f)) & 1) ? 1.0f : 0.35f;
			sigma += s.shell_amp;
			rgb[0] += s.shell_amp * s.shell_col[0] * band; rgb[1] += s.shell_amp * s.shell_col[1] * band; rgb[2] += s.shell_amp * s.shell_col[2] * band;
		}
	}
	if (sigma > 0.f) { const float inv = 1.f / sigma; rgb[0] *= inv; rgb[1] *= inv; rgb[2] *= inv; }
	#pragma unroll
	for (int k = 0; k < 3; ++k) rgb[k] = fminf(fmaxf(rgb[k], 0.f), 1.f);
}

// one thread per pixel; xforms: [n_views] 3x4 column-major camera-to-world in the unit-cube frame
__global__ void __launch_bounds__(256) render_procedural_kernel(Scene scene, uint32_t n_views, int W, int H, float fx, float fy,
                                                                const float* __restrict__ xforms, uint32_t n_steps, uint32_t* __restrict__ out) {
	const uint64_t n_pix = (uint64_t)W * H;
	const uint64_t gid = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
	if (gid >= n_pix * n_views) return;
	const uint32_t view = (uint32_t)(gid / n_pix), pix = (uint32_t)(gid % n_pix);
	const int px = pix % W, py = pix / W;
	const float* xf = xforms + (size_t)view * 12;
	const float cx = (((float)px + 0.5f) / (float)W - 0.5f) * (float)W / fx, cy = (((float)py + 0.5f) / (float)H - 0.5f) * (float)H / fy;
	float dx = xf[0] * cx + xf[3] * cy + xf[6], dy = xf[1] * cx + xf[4] * cy + xf[7], dz = xf[2] * cx + xf[5] * cy + xf[8];
	const float inv = rsqrtf(dx * dx + dy * dy + dz * dz);
	dx *= inv; dy *= inv; dz *= inv;
	const float ox = xf[9], oy = xf[10], oz = xf[11];
	// unit-cube slab test
	float t0 = 0.f, t1 = 1e9f;
	const float o[3] = {ox, oy, oz}, d[3] = {dx, dy, dz};
	#pragma unroll
	for (int k = 0; k < 3; ++k) {
		const float a = (0.f - o[k]) / d[k], b = (1.f - o[k]) / d[k];
		t0 = fmaxf(t0, fminf(a, b)); t1 = fminf(t1, fmaxf(a, b));
	}
	float C[3] = {0.f, 0.f, 0.f}, T = 1.f;
	if (t1 > t0) {
		const float dt = 1.73205080757f / (float)n_steps;
		for (float t = t0 + 0.5f * dt; t < t1 && T > 1e-4f; t += dt) {
			float sigma, rgb[3];
			scene_eval(scene, ox + t * dx, oy + t * dy, oz + t * dz, dx, dy, dz, sigma, rgb);
			if (sigma <= 0.f) continue;
			const float alpha = 1.f - __expf(-sigma * dt), w = alpha * T;
			C[0] += w * rgb[0]; C[1] += w * rgb[1]; C[2] += w * rgb[2];
			T *= 1.f - alpha;
		}
	}
	const float A = 1.f - T;
	uint32_t q[4];
	#pragma unroll
	for (int k = 0; k < 3; ++k) {
		const float straight = A > 1e-6f ? fminf(C[k] / A, 1.f) : 0.f;
		q[k] = (uint32_t)(fminf(fmaxf(linear_to_srgb(straight), 0.f), 1.f) * 255.f + 0.5f);
	}
	q[3] = (uint32_t)(fminf(fmaxf(A, 0.f), 1.f) * 255.f + 0.5f);
	out[gid] = q[0] | (q[1] << 8) | (q[2] << 16) | (q[3] << 24);
}

} // namespace ngp

using namespace ngp;

extern "C" int ngp_render_procedural_views(void* stream, uint32_t seed, uint32_t n_views, int width, int height, float fx, float fy,
                                           const float* d_xforms, uint8_t* d_rgba8) {
	NGP_REQUIRE(d_xforms && d_rgba8, "null pointer");
	NGP_REQUIRE(n_views > 0 && width > 0 && height > 0, "empty image set");
	Scene s;
	Pcg32 rng; rng.seed(seed);
	for (int i = 0; i < NGP_SCENE_BLOBS; ++i) {
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
	const uint64_t total = (uint64_t)n_views * width * height;
	NGP_REQUIRE(total < (1ull << 31) * 256ull, "too many pixels");
	NGP_LAUNCH("render_procedural_kernel", (cudaStream_t)stream, render_procedural_kernel<<<(uint32_t)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(s, n_views, width, height, fx, fy, d_xforms, 1024, (uint32_t*)d_rgba8));
	return check_launch("render_procedural_views");
}
