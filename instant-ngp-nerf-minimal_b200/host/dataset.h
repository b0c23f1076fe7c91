// NeRFSynthetic -- the training set as the reference keeps it (internal/dataset/nerf_synthetic.h:18-41): camera-to-world
// matrices after spec_opengl_to_opencv and all images as linear, alpha-premultiplied fp16 RGBA in device memory.
#pragma once
#include "common.h"

namespace ngp {

struct NeRFSynthetic {
	// transforms_*.json + images (PNG decoded on the host with zlib, or raw .rgba8), nerf_synthetic.cu:31-134
	void load_from_json(const std::string& path, float scale, const float offset[3]);
	// images already in host memory: rgba8 [n][H][W][4] sRGB straight alpha, xforms_ogl [n] 3x4 OpenGL camera-to-world
	void load_from_memory(uint32_t n, int width, int height, float camera_angle_x, const Mat34* xforms_ogl, const uint8_t* rgba8, float scale, const float offset[3]);
	// procedural scene rendered on the device (benchmark fixture; no file I/O): cameras on a sphere, analytic density field
	void load_procedural(uint32_t n, int width, int height, float camera_angle_x, uint32_t seed, float scale, const float offset[3]);

	uint32_t n_images = 0;
	int image_resolution[2] = {0, 0};
	float focal_length[2] = {0, 0};
	float principal_point[2] = {0.5f, 0.5f};
	float scale = 1.f;
	float offset[3] = {0, 0, 0};
	std::vector<std::string> filenames;
	std::vector<Mat34> xforms_host;          // after spec_opengl_to_opencv
	DeviceBuffer<Mat34> xforms;
	DeviceBuffer<precision_t> images;        // [n][H][W][4] fp16
};

// PNG bytes / file -> RGBA8 [h][w][4], converted as stb_image's stbi_load(..., 4) converts (nerf_synthetic.cu:82)
std::vector<uint8_t> decode_png(const std::string& data, int& w, int& h);
std::vector<uint8_t> decode_png_file(const std::string& path, int& w, int& h);

// camera-to-world (OpenGL convention, as transforms_*.json stores it) of a camera on a sphere looking at the origin
Mat34 look_at_origin_opengl(float radius, float theta, float phi);

} // namespace ngp
