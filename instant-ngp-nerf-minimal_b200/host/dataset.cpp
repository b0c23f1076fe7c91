// NeRFSynthetic loaders.  load_from_json follows internal/dataset/nerf_synthetic.cu:31-134 (transforms json, one image
// per frame, spec_opengl_to_opencv on every pose, fov -> focal length, u8 sRGB -> linear premultiplied fp16 on the GPU).
// stb_image is replaced by a PNG reader on top of zlib (every colour type and bit depth, Adam7, tRNS: decode_png below) and a raw
// ".rgba8" fallback; JPEG (stb's other decoder the reference would fall back to) is not implemented.
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

// PNG -> RGBA8 with the conversions stb_image applies for req_comp = 4 (dependencies/stb_image/stb_image.h, stbi__parse_png_file /
// stbi__create_png_image / stbi__convert_format): all colour types (gray, RGB, palette, gray+alpha, RGBA), bit depths 1-16, Adam7
// interlacing, tRNS colour keys and palette alphas; 16-bit samples keep their high byte; sub-byte gray is scaled to 0..255;
// gray -> (g, g, g, a), missing alpha -> 255.  Gamma / colour-profile chunks are ignored, as stb ignores them.
static void png_unfilter(uint8_t* cur, const uint8_t* prev, int filter, size_t stride, size_t bpp) {
	for (size_t x = 0; x < stride; ++x) {
		const int a = x >= bpp ? cur[x - bpp] : 0, b = prev ? prev[x] : 0, c = (prev && x >= bpp) ? prev[x - bpp] : 0;
		int v = cur[x];
		switch (filter) {
			case 0: break;
			case 1: v += a; break;
			case 2: v += b; break;
			case 3: v += (a + b) >> 1; break;
			case 4: { const int pa = abs(b - c), pb = abs(a - c), pc = abs(a + b - 2 * c); v += (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); break; }
			default: throw std::runtime_error("PNG: bad filter type");
		}
		cur[x] = (uint8_t)v;
	}
}

std::vector<uint8_t> decode_png(const std::string& data, int& w, int& h) {
	static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
	if (data.size() < 8 || memcmp(data.data(), sig, 8) != 0) throw std::runtime_error("not a PNG file");
	const uint8_t* p = (const uint8_t*)data.data();
	size_t pos = 8;
	int depth = 0, ctype = 0, interlace = 0;
	std::string idat;
	uint8_t palette[256][4];
	for (auto& e : palette) { e[0] = e[1] = e[2] = 0; e[3] = 255; }
	bool has_key = false; uint16_t key[3] = {0, 0, 0};
	bool seen_ihdr = false;
	while (pos + 12 <= data.size()) {
		const uint32_t len = be32(p + pos);
		const std::string type(data, pos + 4, 4);
		if (pos + 12 + (size_t)len > data.size()) throw std::runtime_error("PNG: truncated chunk");
		const uint8_t* body = p + pos + 8;
		if (type == "IHDR") {
			if (len < 13) throw std::runtime_error("PNG: bad IHDR");
			w = (int)be32(body); h = (int)be32(body + 4); depth = body[8]; ctype = body[9]; interlace = body[12]; seen_ihdr = true;
		} else if (type == "PLTE") {
			for (uint32_t k = 0; k < len / 3 && k < 256; ++k) { palette[k][0] = body[3 * k]; palette[k][1] = body[3 * k + 1]; palette[k][2] = body[3 * k + 2]; }
		} else if (type == "tRNS") {
			if (ctype == 3) { for (uint32_t k = 0; k < len && k < 256; ++k) palette[k][3] = body[k]; }
			else if (ctype == 0 && len >= 2) { has_key = true; key[0] = (uint16_t)((body[0] << 8) | body[1]); }
			else if (ctype == 2 && len >= 6) { has_key = true; for (int k = 0; k < 3; ++k) key[k] = (uint16_t)((body[2 * k] << 8) | body[2 * k + 1]); }
		} else if (type == "IDAT") idat.append((const char*)body, len);
		else if (type == "IEND") break;
		pos += 12 + (size_t)len;
	}
	if (!seen_ihdr || w <= 0 || h <= 0) throw std::runtime_error("PNG: missing IHDR");
	const bool depth_ok = depth == 8 || depth == 16 || ((ctype == 0 || ctype == 3) && (depth == 1 || depth == 2 || depth == 4));
	if (!depth_ok || (ctype == 3 && depth == 16) || interlace > 1 || !(ctype == 0 || ctype == 2 || ctype == 3 || ctype == 4 || ctype == 6))
		throw std::runtime_error("PNG: unsupported colour type / bit depth");
	const int ch = ctype == 0 || ctype == 3 ? 1 : ctype == 2 ? 3 : ctype == 4 ? 2 : 4;
	const size_t bits_pp = (size_t)ch * depth, bpp = std::max<size_t>(1, bits_pp / 8);
	auto row_bytes = [&](int pw) { return ((size_t)pw * bits_pp + 7) / 8; };

	// passes: one for a plain image, seven for Adam7
	struct Pass { int x0, y0, dx, dy; };
	static const Pass adam7[7] = {{0, 0, 8, 8}, {4, 0, 8, 8}, {0, 4, 4, 8}, {2, 0, 4, 4}, {0, 2, 2, 4}, {1, 0, 2, 2}, {0, 1, 1, 2}};
	const Pass plain = {0, 0, 1, 1};
	size_t raw_size = 0;
	for (int k = 0; k < (interlace ? 7 : 1); ++k) {
		const Pass& ps = interlace ? adam7[k] : plain;
		const int pw = (w - ps.x0 + ps.dx - 1) / ps.dx, ph = (h - ps.y0 + ps.dy - 1) / ps.dy;
		if (pw > 0 && ph > 0) raw_size += (row_bytes(pw) + 1) * ph;
	}
	std::vector<uint8_t> raw(raw_size);
	uLongf out_len = (uLongf)raw.size();
	if (uncompress(raw.data(), &out_len, (const Bytef*)idat.data(), (uLong)idat.size()) != Z_OK || out_len != raw.size())
		throw std::runtime_error("PNG: inflate failed");

	// one sample of the unfiltered row as a 16-bit value (depth < 16: the raw sample)
	auto sample = [&](const uint8_t* row, size_t px, int c) -> uint16_t {
		if (depth == 8) return row[px * ch + c];
		if (depth == 16) return (uint16_t)((row[(px * ch + c) * 2] << 8) | row[(px * ch + c) * 2 + 1]);
		const size_t bit = px * depth;
		return (uint16_t)((row[bit >> 3] >> (8 - depth - (bit & 7))) & ((1u << depth) - 1));
	};
	static const int depth_scale[9] = {0, 0xff, 0x55, 0, 0x11, 0, 0, 0, 0x01};
	std::vector<uint8_t> rgba((size_t)w * h * 4);
	size_t off = 0;
	for (int k = 0; k < (interlace ? 7 : 1); ++k) {
		const Pass& ps = interlace ? adam7[k] : plain;
		const int pw = (w - ps.x0 + ps.dx - 1) / ps.dx, ph = (h - ps.y0 + ps.dy - 1) / ps.dy;
		if (pw <= 0 || ph <= 0) continue;
		const size_t stride = row_bytes(pw);
		const uint8_t* prev = nullptr;
		for (int y = 0; y < ph; ++y) {
			uint8_t* row = raw.data() + off;
			const int filter = row[0]; ++row;
			png_unfilter(row, prev, filter, stride, bpp);
			prev = row;
			off += stride + 1;
			for (int x = 0; x < pw; ++x) {
				uint8_t* d = rgba.data() + ((size_t)(ps.y0 + y * ps.dy) * w + (ps.x0 + x * ps.dx)) * 4;
				auto to8 = [&](uint16_t v) -> uint8_t { return depth == 16 ? (uint8_t)(v >> 8) : depth == 8 ? (uint8_t)v : (uint8_t)(v * depth_scale[depth]); };
				if (ctype == 3) { const uint8_t* e = palette[sample(row, x, 0) & 255]; d[0] = e[0]; d[1] = e[1]; d[2] = e[2]; d[3] = e[3]; }
				else if (ctype == 0) {
					const uint16_t g = sample(row, x, 0);
					d[0] = d[1] = d[2] = to8(g);
					// stb compares the colour key after scaling sub-byte gray to 8 bits, 16-bit samples in full
					d[3] = has_key && (depth == 16 ? g == key[0] : to8(g) == (uint8_t)((key[0] & 255) * (depth < 8 ? depth_scale[depth] : 1))) ? 0 : 255;
				} else if (ctype == 4) { d[0] = d[1] = d[2] = to8(sample(row, x, 0)); d[3] = to8(sample(row, x, 1)); }
				else {
					const uint16_t r = sample(row, x, 0), g = sample(row, x, 1), b = sample(row, x, 2);
					d[0] = to8(r); d[1] = to8(g); d[2] = to8(b);
					if (ctype == 6) d[3] = to8(sample(row, x, 3));
					else d[3] = has_key && (depth == 16 ? (r == key[0] && g == key[1] && b == key[2]) : (d[0] == (key[0] & 255) && d[1] == (key[1] & 255) && d[2] == (key[2] & 255))) ? 0 : 255;
				}
			}
		}
	}
	return rgba;
}

std::vector<uint8_t> decode_png_file(const std::string& path, int& w, int& h) { return decode_png(read_file(path), w, h); }

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
