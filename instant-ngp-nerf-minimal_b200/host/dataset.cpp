// NeRFSynthetic loaders.  load_from_json follows internal/dataset/nerf_synthetic.cu:31-134 (transforms json, one image
// per frame, spec_opengl_to_opencv on every pose, fov -> focal length, u8 sRGB -> linear premultiplied fp16 on the GPU).
// stb_image is replaced by a small PNG reader on top of zlib (8-bit gray / gray+alpha / RGB / RGBA, non-interlaced) and a
// raw ".rgba8" fallback.
#include "dataset.h"
#include "json.h"
#include <zlib.h>
#include <fstream>
#include <sstream>
#include <sys/stat.h>

namespace ngp {

static bool file_exists(const std::string& p) { struct stat st; return stat(p.c_str(), &st) == 0; }
static std::string read_file(const std::string& p) {
	std::ifstream f(p, std::ios::in | std::ios::binary);
	if (!f) throw std::runtime_error("cannot open " + p);
	std::stringstream ss; ss << f.rdbuf();
	return ss.str();
}
static std::string parent_path(const std::string& p) { const size_t k = p.find_last_of('/'); return k == std::string::npos ? "." : p.substr(0, k); }
static std::string extension(const std::string& p) {
	const size_t s = p.find_last_of('/'), k = p.find_last_of('.');
	return (k == std::string::npos || (s != std::string::npos && k < s)) ? "" : p.substr(k + 1);
}

static uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

// decode to RGBA8; throws on anything but 8-bit non-interlaced colour types 0/2/4/6
static std::vector<uint8_t> decode_png(const std::string& data, int& w, int& h) {
	static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
	if (data.size() < 8 || memcmp(data.data(), sig, 8) != 0) throw std::runtime_error("not a PNG file");
	const uint8_t* p = (const uint8_t*)data.data();
	size_t pos = 8;
	int depth = 0, ctype = 0, interlace = 0;
	std::string idat;
	while (pos + 12 <= data.size()) {
		const uint32_t len = be32(p + pos);
		const std::string type(data, pos + 4, 4);
		if (pos + 12 + len > data.size()) throw std::runtime_error("PNG: truncated chunk");
		const uint8_t* body = p + pos + 8;
		if (type == "IHDR") { w = (int)be32(body); h = (int)be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12]; }
		else if (type == "IDAT") idat.append((const char*)body, len);
		else if (type == "IEND") break;
		pos += 12 + len;
	}
	if (depth != 8 || interlace != 0 || !(ctype == 0 || ctype == 2 || ctype == 4 || ctype == 6))
		throw std::runtime_error("PNG: only 8-bit non-interlaced gray/RGB(A) images are supported");
	const int ch = ctype == 0 ? 1 : ctype == 2 ? 3 : ctype == 4 ? 2 : 4;
	const size_t stride = (size_t)w * ch;
	std::vector<uint8_t> raw((stride + 1) * h);
	uLongf out_len = (uLongf)raw.size();
	if (uncompress(raw.data(), &out_len, (const Bytef*)idat.data(), (uLong)idat.size()) != Z_OK || out_len != raw.size())
		throw std::runtime_error("PNG: inflate failed");
	std::vector<uint8_t> img(stride * h);
	for (int y = 0; y < h; ++y) {
		const uint8_t* src = raw.data() + (stride + 1) * y;
		const int filter = src[0]; ++src;
		uint8_t* dst = img.data() + stride * y;
		const uint8_t* up = y ? dst - stride : nullptr;
		for (size_t x = 0; x < stride; ++x) {
			const int a = x >= (size_t)ch ? dst[x - ch] : 0, b = up ? up[x] : 0, c = (up && x >= (size_t)ch) ? up[x - ch] : 0;
			int v = src[x];
			switch (filter) {
				case 0: break;
				case 1: v += a; break;
				case 2: v += b; break;
				case 3: v += (a + b) >> 1; break;
				case 4: { const int pa = abs(b - c), pb = abs(a - c), pc = abs(a + b - 2 * c); v += (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
				default: throw std::runtime_error("PNG: bad filter type");
			}
			dst[x] = (uint8_t)v;
		}
	}
	std::vector<uint8_t> rgba((size_t)w * h * 4);
	for (size_t i = 0; i < (size_t)w * h; ++i) {
		const uint8_t* s = img.data() + i * ch;
		uint8_t* d = rgba.data() + i * 4;
		if (ch == 1) { d[0] = d[1] = d[2] = s[0]; d[3] = 255; }
		else if (ch == 2) { d[0] = d[1] = d[2] = s[0]; d[3] = s[1]; }
		else if (ch == 3) { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; d[3] = 255; }
		else { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; d[3] = s[3]; }
	}
	return rgba;
}

void NeRFSynthetic::load_from_memory(uint32_t n, int width, int height, float camera_angle_x, const Mat34* xforms_ogl, const uint8_t* rgba8, float scale_, const float offset_[3]) {
	n_images = n; image_resolution[0] = width; image_resolution[1] = height;
	scale = scale_; for (int k = 0; k < 3; ++k) offset[k] = offset_[k];
	xforms_host.resize(n);
	for (uint32_t i = 0; i < n; ++i) xforms_host[i] = spec_opengl_to_opencv(xforms_ogl[i], scale, offset);
	focal_length[0] = focal_length[1] = fov_to_focal_length(width, camera_angle_x);
	const size_t n_pixels = (size_t)width * height * n;
	DeviceBuffer<uint8_t> staging(n_pixels * 4);
	NGP_CUDA_CHECK(cudaMemcpy(staging.data(), rgba8, n_pixels * 4, cudaMemcpyHostToDevice));
	images.resize(n_pixels * 4);
	api_check(ngp_from_rgba32(nullptr, n_pixels, staging.data(), images.data()));
	xforms.resize_and_copy_from_host(xforms_host);
	NGP_CUDA_CHECK(cudaDeviceSynchronize());
}

void NeRFSynthetic::load_from_json(const std::string& path, float scale_, const float offset_[3]) {
	const std::string base = parent_path(path);
	const Json json = Json::parse(read_file(path));
	const Json& frames = json.at("frames");
	const uint32_t n = (uint32_t)frames.size();
	if (!n) throw std::runtime_error("dataset has no frames: " + path);
	std::vector<Mat34> ogl(n);
	std::vector<uint8_t> all;
	int W = 0, H = 0;
	filenames.resize(n);
	for (uint32_t i = 0; i < n; ++i) {
		const Json& frame = frames[i];
		std::string p = base + "/" + frame.at("file_path").str();
		if (extension(p).empty()) {
			if (file_exists(p + ".png")) p += ".png";
			else if (file_exists(p + ".rgba8")) p += ".rgba8";
			else throw std::runtime_error("Could not find image file: " + p + ".png");
		}
		filenames[i] = p.substr(p.find_last_of('/') + 1);
		int w = 0, h = 0;
		std::vector<uint8_t> img;
		if (extension(p) == "rgba8") {   // raw: two little-endian int32 (w, h) then w*h*4 bytes
			const std::string d = read_file(p);
			if (d.size() < 8) throw std::runtime_error("image not found: " + p);
			memcpy(&w, d.data(), 4); memcpy(&h, d.data() + 4, 4);
			if (d.size() != 8 + (size_t)w * h * 4) throw std::runtime_error("bad .rgba8 file: " + p);
			img.assign(d.begin() + 8, d.end());
		} else if (extension(p) == "png") {
			img = decode_png(read_file(p), w, h);
		} else throw std::runtime_error("unsupported image type (PNG or .rgba8): " + p);
		if (W && (w != W || h != H)) throw std::runtime_error("training images are not all the same size");
		W = w; H = h;
		all.insert(all.end(), img.begin(), img.end());
		const Json& tm = frame.at("transform_matrix");
		for (int m = 0; m < 3; ++m) for (int k = 0; k < 4; ++k) ogl[i].at(m, k) = (float)tm[m][k].number();
	}
	load_from_memory(n, W, H, (float)json.at("camera_angle_x").number(), ogl.data(), all.data(), scale_, offset_);
}

Mat34 look_at_origin_opengl(float radius, float theta, float phi) {
	// camera position on the sphere (Blender: z up); OpenGL camera looks down -z, +y up
	const float cx = radius * cosf(phi) * cosf(theta), cy = radius * cosf(phi) * sinf(theta), cz = radius * sinf(phi);
	float f[3] = {-cx, -cy, -cz};
	const float fn = sqrtf(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
	for (float& v : f) v /= fn;
	const float up[3] = {0.f, 0.f, 1.f};
	float r[3] = {f[1] * up[2] - f[2] * up[1], f[2] * up[0] - f[0] * up[2], f[0] * up[1] - f[1] * up[0]};
	const float rn = sqrtf(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
	for (float& v : r) v /= rn;
	const float u[3] = {r[1] * f[2] - r[2] * f[1], r[2] * f[0] - r[0] * f[2], r[0] * f[1] - r[1] * f[0]};
	Mat34 m;
	for (int k = 0; k < 3; ++k) { m.at(k, 0) = r[k]; m.at(k, 1) = u[k]; m.at(k, 2) = -f[k]; }
	m.at(0, 3) = cx; m.at(1, 3) = cy; m.at(2, 3) = cz;
	return m;
}

void NeRFSynthetic::load_procedural(uint32_t n, int width, int height, float camera_angle_x, uint32_t seed, float scale_, const float offset_[3]) {
	n_images = n; image_resolution[0] = width; image_resolution[1] = height;
	scale = scale_; for (int k = 0; k < 3; ++k) offset[k] = offset_[k];
	// cameras on the upper hemisphere of radius 4.0311 (the Blender nerf_synthetic rig), golden-angle spiral
	Pcg32 rng{(uint64_t)seed};
	xforms_host.resize(n);
	filenames.resize(n);
	for (uint32_t i = 0; i < n; ++i) {
		const float theta = 2.39996323f * (float)i + 0.1f * rng.next_float();
		const float phi = asinf(0.05f + 0.9f * ((float)i + 0.5f) / (float)n);
		xforms_host[i] = spec_opengl_to_opencv(look_at_origin_opengl(4.0311f, theta, phi), scale, offset);
		filenames[i] = "procedural_" + std::to_string(i);
	}
	focal_length[0] = focal_length[1] = fov_to_focal_length(width, camera_angle_x);
	xforms.resize_and_copy_from_host(xforms_host);
	const size_t n_pixels = (size_t)width * height * n;
	DeviceBuffer<uint8_t> staging(n_pixels * 4);
	api_check(ngp_render_procedural_views(nullptr, seed, n, width, height, focal_length[0], focal_length[1], (const float*)xforms.data(), staging.data()));
	images.resize(n_pixels * 4);
	api_check(ngp_from_rgba32(nullptr, n_pixels, staging.data(), images.data()));
	NGP_CUDA_CHECK(cudaDeviceSynchronize());
}

} // namespace ngp
