// TEST INFRASTRUCTURE ONLY (see oracle_common.h).  Pinned to golden vectors from the reference's CUDA build (tests/golden/).
// CPU restatement of the NeRF network dataflow:
//   spherical harmonics degree 4 : tiny-cuda-nn/encodings/spherical_harmonics.h:63-100
//   fully fused MLPs             : tiny-cuda-nn/src/fully_fused_mlp.cu:480-533 (fwd), :152-248 (bwd),
//                                  weight layout :647-665 (row-major [out][in], layers contiguous)
//   dataflow / glue              : internal/network/ngp_network.h:44-127, 217-494
// Weights arena order: W0[64][32] W1[16][64] | V0[64][32] V1[64][64] V2[16][64]  (10240 halfs).
#include "oracle_common.h"
#include <omp.h>

using namespace orc;

namespace {

constexpr int W0 = 0, W1 = 2048, V0 = 3072, V1 = 5120, V2 = 9216, N_MLP = 10240;

// y[out] = W[out][in] . x[in]; accumulate mode 0: fp32 running sum, one fp16 rounding at the end.
// mode 1: emulate fp16 accumulators (k-blocks of 16 summed in fp32, then added into an fp16 accumulator)
static inline void matvec(const float* W, int n_out, int n_in, const float* x, float* y, bool relu, int mode) {
	for (int o = 0; o < n_out; ++o) {
		const float* w = W + (size_t)o * n_in;
		float r;
		if (mode == 0) {
			float acc = 0.f;
			for (int k = 0; k < n_in; ++k) acc = fmaf(w[k], x[k], acc);
			r = h16r(acc);
		} else {
			float acc = 0.f;
			for (int kb = 0; kb < n_in; kb += 16) {
				float part = 0.f;
				for (int k = kb; k < kb + 16; ++k) part = fmaf(w[k], x[k], part);
				acc = h16r(acc + part);
			}
			r = acc;
		}
		if (relu) r = r > 0.f ? r : 0.f;
		y[o] = r;
	}
}

// x[in] = W^T[in][out] . g[out]
static inline void matvec_t(const float* W, int n_out, int n_in, const float* g, float* x) {
	for (int k = 0; k < n_in; ++k) {
		float acc = 0.f;
		for (int o = 0; o < n_out; ++o) acc = fmaf(W[(size_t)o * n_in + k], g[o], acc);
		x[k] = h16r(acc);
	}
}

static inline void sh4(float dx, float dy, float dz, float* out) {
	float x = dx * 2.f - 1.f, y = dy * 2.f - 1.f, z = dz * 2.f - 1.f;
	float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
	out[0] = 0.28209479177387814f;
	out[1] = -0.48860251190291987f * y;
	out[2] = 0.48860251190291987f * z;
	out[3] = -0.48860251190291987f * x;
	out[4] = 1.0925484305920792f * xy;
	out[5] = -1.0925484305920792f * yz;
	out[6] = 0.94617469575755997f * z2 - 0.31539156525251999f;
	out[7] = -1.0925484305920792f * xz;
	out[8] = 0.54627421529603959f * x2 - 0.54627421529603959f * y2;
	out[9] = 0.59004358992664352f * y * (-3.0f * x2 + y2);
	out[10] = 2.8906114426405538f * xy * z;
	out[11] = 0.45704579946446572f * y * (1.0f - 5.0f * z2);
	out[12] = 0.3731763325901154f * z * (5.0f * z2 - 3.0f);
	out[13] = 0.45704579946446572f * x * (1.0f - 5.0f * z2);
	out[14] = 1.4453057213202769f * z * (x2 - y2);
	out[15] = 0.59004358992664352f * x * (-x2 + 3.0f * y2);
	for (int i = 0; i < 16; ++i) out[i] = h16r(out[i]);
}

struct Acts { float e[32], h[64], o[16], x[32], h1[64], h2[64], c[16]; };

static inline void forward_one(const float* Wf, const uint16_t* enc_row, const float* dir, Acts& a, int mode) {
	for (int k = 0; k < 32; ++k) a.e[k] = h16from(enc_row[k]);
	matvec(Wf + W0, 64, 32, a.e, a.h, true, mode);
	matvec(Wf + W1, 16, 64, a.h, a.o, false, mode);
	for (int k = 0; k < 16; ++k) a.x[k] = a.o[k];
	if (dir) sh4(dir[0], dir[1], dir[2], a.x + 16);
	else for (int k = 16; k < 32; ++k) a.x[k] = 0.f;
	matvec(Wf + V0, 64, 32, a.x, a.h1, true, mode);
	matvec(Wf + V1, 64, 64, a.h1, a.h2, true, mode);
	matvec(Wf + V2, 16, 64, a.h2, a.c, false, mode);
}

} // namespace

extern "C" {

void orc_sh4(const float* dirs, uint32_t stride_floats, uint32_t n, uint16_t* out /*[n][16]*/) {
	for (uint32_t i = 0; i < n; ++i) {
		float o[16];
		sh4(dirs[(size_t)i * stride_floats], dirs[(size_t)i * stride_floats + 1], dirs[(size_t)i * stride_floats + 2], o);
		for (int k = 0; k < 16; ++k) out[(size_t)i * 16 + k] = h16bits(o[k]);
	}
}

// density only: out [n][16] fp16 (reference NGPNetwork::density, ngp_network.h:286-307)
void orc_density_fwd(const uint16_t* weights, const uint16_t* enc, uint32_t n, int mode, uint16_t* out) {
	std::vector<float> Wf(N_MLP);
	for (int i = 0; i < N_MLP; ++i) Wf[i] = h16from(weights[i]);
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)n; ++i) {
		float e[32], h[64], o[16];
		for (int k = 0; k < 32; ++k) e[k] = h16from(enc[(size_t)i * 32 + k]);
		matvec(Wf.data() + W0, 64, 32, e, h, true, mode);
		matvec(Wf.data() + W1, 16, 64, h, o, false, mode);
		for (int k = 0; k < 16; ++k) out[(size_t)i * 16 + k] = h16bits(o[k]);
	}
}

// full network forward: enc [n][32] fp16, coords (Coordinate AoS, dir at float offset 4) -> rgbsigma [n][4] fp16
void orc_network_fwd(const uint16_t* weights, const uint16_t* enc, const float* coords, uint32_t coord_stride,
                     uint32_t n, int mode, uint16_t* rgbsigma) {
	std::vector<float> Wf(N_MLP);
	for (int i = 0; i < N_MLP; ++i) Wf[i] = h16from(weights[i]);
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)n; ++i) {
		Acts a;
		forward_one(Wf.data(), enc + (size_t)i * 32, coords + (size_t)i * coord_stride + 4, a, mode);
		rgbsigma[(size_t)i * 4 + 0] = h16bits(a.c[0]);
		rgbsigma[(size_t)i * 4 + 1] = h16bits(a.c[1]);
		rgbsigma[(size_t)i * 4 + 2] = h16bits(a.c[2]);
		rgbsigma[(size_t)i * 4 + 3] = h16bits(a.o[0]);
	}
}

// forward + backward over a batch.  dL: [n][4] fp16 (r,g,b,sigma).  Outputs: wgrad [10240] f32 (exact
// sums of fp32 products, deterministic order), denc [n][32] fp16.
void orc_network_fwd_bwd(const uint16_t* weights, const uint16_t* enc, const float* coords, uint32_t coord_stride,
                         const uint16_t* dL, uint32_t n, int mode, float* wgrad, uint16_t* denc, uint16_t* rgbsigma) {
	std::vector<float> Wf(N_MLP);
	for (int i = 0; i < N_MLP; ++i) Wf[i] = h16from(weights[i]);
	const int nt = omp_get_max_threads();
	std::vector<std::vector<double>> part(nt, std::vector<double>(N_MLP, 0.0));
	#pragma omp parallel
	{
		std::vector<double>& G = part[omp_get_thread_num()];
		#pragma omp for schedule(static)
		for (int64_t i = 0; i < (int64_t)n; ++i) {
			Acts a;
			forward_one(Wf.data(), enc + (size_t)i * 32, coords + (size_t)i * coord_stride + 4, a, mode);
			if (rgbsigma) {
				for (int k = 0; k < 3; ++k) rgbsigma[(size_t)i * 4 + k] = h16bits(a.c[k]);
				rgbsigma[(size_t)i * 4 + 3] = h16bits(a.o[0]);
			}
			float dc[16] = {0}, dh2[64], dh1[64], dx[32], dout[16], dh[64], de[32];
			for (int k = 0; k < 3; ++k) dc[k] = h16from(dL[(size_t)i * 4 + k]);
			const float dsigma = h16from(dL[(size_t)i * 4 + 3]);
			for (int o = 0; o < 16; ++o) for (int k = 0; k < 64; ++k) G[V2 + o * 64 + k] += (double)dc[o] * a.h2[k];
			matvec_t(Wf.data() + V2, 16, 64, dc, dh2);
			for (int k = 0; k < 64; ++k) if (!(a.h2[k] > 0.f)) dh2[k] = 0.f;
			for (int o = 0; o < 64; ++o) for (int k = 0; k < 64; ++k) G[V1 + o * 64 + k] += (double)dh2[o] * a.h1[k];
			matvec_t(Wf.data() + V1, 64, 64, dh2, dh1);
			for (int k = 0; k < 64; ++k) if (!(a.h1[k] > 0.f)) dh1[k] = 0.f;
			for (int o = 0; o < 64; ++o) for (int k = 0; k < 32; ++k) G[V0 + o * 32 + k] += (double)dh1[o] * a.x[k];
			matvec_t(Wf.data() + V0, 64, 32, dh1, dx);
			for (int k = 0; k < 16; ++k) dout[k] = h16r(dx[k] + (k == 0 ? dsigma : 0.f));
			for (int o = 0; o < 16; ++o) for (int k = 0; k < 64; ++k) G[W1 + o * 64 + k] += (double)dout[o] * a.h[k];
			matvec_t(Wf.data() + W1, 16, 64, dout, dh);
			for (int k = 0; k < 64; ++k) if (!(a.h[k] > 0.f)) dh[k] = 0.f;
			for (int o = 0; o < 64; ++o) for (int k = 0; k < 32; ++k) G[W0 + o * 32 + k] += (double)dh[o] * a.e[k];
			matvec_t(Wf.data() + W0, 64, 32, dh, de);
			if (denc) for (int k = 0; k < 32; ++k) denc[(size_t)i * 32 + k] = h16bits(de[k]);
		}
	}
	for (int k = 0; k < N_MLP; ++k) {
		double s = 0.0;
		for (int t = 0; t < nt; ++t) s += part[t][k];
		wgrad[k] = (float)s;
	}
}

} // extern "C"
