// CudaRenderBuffer -- headless stand-in for internal/utils/render_buffer.{h,cu}: the reference accumulates into a float4
// frame buffer and tonemaps into a cudaArray surface; here the surface is a linear float4 device buffer (no graphics
// interop exists on this box), everything else (running spp mean, exposure, background blend, tonemap, sRGB) is the same.
#pragma once
#include "common.h"

namespace ngp {

class CudaRenderBuffer {
public:
	void resize(int w, int h) {
		if (w == m_res[0] && h == m_res[1]) return;
		m_res[0] = w; m_res[1] = h;
		const size_t n = (size_t)w * h * 4;
		m_frame.resize(n); m_accumulate.resize(n); m_surface.resize(n);
		m_row0 = 0; m_row1 = h;
		reset_accumulation();
	}
	const int* resolution() const { return m_res; }
	void reset_accumulation() { m_spp = 0; }
	uint32_t spp() const { return m_spp; }
	// tile-sharded rendering: clear / accumulate / tonemap touch the rows [r0, r1) only (r1 < 0: the whole frame)
	void set_rows(int r0, int r1) { m_row0 = r1 < 0 ? 0 : r0; m_row1 = r1 < 0 ? m_res[1] : r1; }
	size_t first_pixel() const { return (size_t)m_row0 * m_res[0]; }
	uint32_t n_band_pixels() const { return (uint32_t)((m_row1 - m_row0) * m_res[0]); }
	void clear_frame_buffer(cudaStream_t s) {
		if (n_band_pixels()) NGP_CUDA_CHECK(cudaMemsetAsync(m_frame.data() + first_pixel() * 4, 0, (size_t)n_band_pixels() * 4 * sizeof(float), s));
	}
	float* frame_buffer() const { return m_frame.data(); }
	float* surface() const { return m_surface.data(); }
	void set_color_space(EColorSpace c) { m_color_space = c; }
	void set_tonemap_curve(ETonemapCurve c) { m_curve = c; }
	void accumulate(cudaStream_t s) {
		const uint32_t n = n_band_pixels();
		const size_t o = first_pixel() * 4;
		if (n && m_spp == 0) NGP_CUDA_CHECK(cudaMemsetAsync(m_accumulate.data() + o, 0, (size_t)n * 4 * sizeof(float), s));
		if (n) api_check(ngp_accumulate(s, n, m_frame.data() + o, m_accumulate.data() + o, (float)m_spp, (int)m_color_space));
		++m_spp;
	}
	void tonemap(float exposure, const float background_color[4], EColorSpace output_color_space, cudaStream_t s) {
		const uint32_t n = n_band_pixels();
		const size_t o = first_pixel() * 4;
		if (n) api_check(ngp_tonemap(s, n, exposure, background_color, m_accumulate.data() + o, (int)m_color_space, (int)output_color_space, (int)m_curve, m_surface.data() + o));
	}
private:
	int m_res[2] = {0, 0};
	int m_row0 = 0, m_row1 = 0;
	uint32_t m_spp = 0;
	EColorSpace m_color_space = EColorSpace::Linear;
	ETonemapCurve m_curve = ETonemapCurve::Identity;
	DeviceBuffer<float> m_frame, m_accumulate, m_surface;
};

} // namespace ngp
