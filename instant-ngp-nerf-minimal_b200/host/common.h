// Host-side types shared by the sampler / network / render / testbed classes.
// Mirrors internal/utils/common.h:47-153 of the reference (constants, enums, wire formats) and the two tiny-cuda-nn
// utilities the host code leans on: pcg32 (dependencies/pcg32/pcg32.h:44-205) and GPUMemory (gpu_memory.h).
#pragma once
#include <cuda_runtime_api.h>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <random>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/ngp_b200.h"

namespace ngp {

static constexpr uint32_t MARCH_ITER = 10000;            // common.h:57
static constexpr uint32_t BATCH_SIZE_MULTIPLE = 256;     // common.h:59
static constexpr float LOSS_SCALE = 128.f;               // common.h:61

enum class EActivation : int { None = 0, ReLU, Logistic, Exponential };
enum class ERenderMode : int { AO = 0, Shade, Normals, Depth, Distance, Stepsize, Cost };
enum class EColorSpace : int { Linear = 0, SRGB };
enum class ETonemapCurve : int { Identity = 0, ACES, Hable, Reinhard };
enum class ELossType : int { L2 = 0, L1, Mape, Smape, SmoothL1, LogL1, RelativeL2 };

inline ELossType string_to_loss_type(const std::string& s0) {
	std::string s; for (char c : s0) s.push_back((char)tolower(c));
	if (s == "l2") return ELossType::L2;
	if (s == "l1") return ELossType::L1;
	if (s == "mape") return ELossType::Mape;
	if (s == "smape") return ELossType::Smape;
	if (s == "smoothl1" || s == "huber") return ELossType::SmoothL1;
	if (s == "logl1") return ELossType::LogL1;
	if (s == "relativel2") return ELossType::RelativeL2;
	throw std::runtime_error("Unknown loss type: " + s0);
}

struct Ray { float o[3], d[3]; };                                  // 24 B
struct Position { float p[3]; };                                   // 12 B
struct Coordinate { float pos[3]; float dt; float dir[3]; };       // 28 B
struct RayPayload { uint32_t alive; float o[3], d[3]; uint32_t idx; float t; uint32_t n_steps; };  // 40 B
static_assert(sizeof(Coordinate) == 28 && sizeof(RayPayload) == 40 && sizeof(Ray) == 24, "wire formats");

inline uint32_t next_multiple(uint32_t v, uint32_t d) { return ((v + d - 1) / d) * d; }

inline void cuda_check(cudaError_t e, const char* what) {
	if (e != cudaSuccess) throw std::runtime_error(std::string(what) + " failed: " + cudaGetErrorString(e));
}
#define NGP_CUDA_CHECK(x) ::ngp::cuda_check((x), #x)

// status of a C-ABI call -> exception (the reference surfaces std::runtime_error through pybind)
inline void api_check(int rc) { if (rc != 0) throw std::runtime_error(ngp_last_error()); }

// ---- pcg32, the generator behind tcnn::default_rng_t -----------------------------------------------
struct Pcg32 {
	uint64_t state = 0x853c49e6748fea9bULL, inc = 0xda3e39cb94b95bdbULL;
	Pcg32() {}
	explicit Pcg32(uint64_t initstate, uint64_t initseq = 1u) { seed(initstate, initseq); }
	void seed(uint64_t initstate, uint64_t initseq = 1u) { state = 0u; inc = (initseq << 1u) | 1u; next_uint(); state += initstate; next_uint(); }
	uint32_t next_uint() {
		const uint64_t old = state;
		state = old * 0x5851f42d4c957f2dULL + inc;
		const uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u), rot = (uint32_t)(old >> 59u);
		return (xs >> rot) | (xs << ((~rot + 1u) & 31));
	}
	float next_float() { const uint32_t u = (next_uint() >> 9) | 0x3f800000u; float f; std::memcpy(&f, &u, 4); return f - 1.0f; }
	void advance(int64_t delta_ = (1ll << 32)) {
		uint64_t cur_mult = 0x5851f42d4c957f2dULL, cur_plus = inc, acc_mult = 1u, acc_plus = 0u, delta = (uint64_t)delta_;
		while (delta > 0) {
			if (delta & 1) { acc_mult *= cur_mult; acc_plus = acc_plus * cur_mult + cur_plus; }
			cur_plus = (cur_mult + 1) * cur_plus; cur_mult *= cur_mult; delta /= 2;
		}
		state = acc_mult * state + acc_plus;
	}
};
using default_rng_t = Pcg32;

// ---- device memory, GPUMemory<T>-like: grows, never shrinks, freed with the owner --------------------
template <typename T>
class DeviceBuffer {
public:
	DeviceBuffer() {}
	explicit DeviceBuffer(size_t n) { resize(n); }
	DeviceBuffer(const DeviceBuffer&) = delete;
	DeviceBuffer& operator=(const DeviceBuffer&) = delete;
	~DeviceBuffer() { if (m_ptr) cudaFree(m_ptr); }
	void resize(size_t n) {
		if (n == m_size) return;
		if (m_ptr) { cudaFree(m_ptr); m_ptr = nullptr; }
		m_size = n;
		if (n) NGP_CUDA_CHECK(cudaMalloc((void**)&m_ptr, n * sizeof(T)));
	}
	void enlarge(size_t n) { if (n > m_size) resize(n); }
	void memset(int v, cudaStream_t s = nullptr) { if (m_size) NGP_CUDA_CHECK(cudaMemsetAsync(m_ptr, v, m_size * sizeof(T), s)); }
	void copy_from_host(const T* src, size_t n, cudaStream_t s = nullptr) {
		NGP_CUDA_CHECK(cudaMemcpyAsync(m_ptr, src, n * sizeof(T), cudaMemcpyHostToDevice, s));
		NGP_CUDA_CHECK(cudaStreamSynchronize(s));
	}
	void resize_and_copy_from_host(const std::vector<T>& v) { resize(v.size()); if (!v.empty()) copy_from_host(v.data(), v.size()); }
	void copy_to_host(T* dst, size_t n, cudaStream_t s = nullptr) const {
		NGP_CUDA_CHECK(cudaMemcpyAsync(dst, m_ptr, n * sizeof(T), cudaMemcpyDeviceToHost, s));
		NGP_CUDA_CHECK(cudaStreamSynchronize(s));
	}
	T* data() const { return m_ptr; }
	size_t size() const { return m_size; }
	size_t get_bytes() const { return m_size * sizeof(T); }
private:
	T* m_ptr = nullptr;
	size_t m_size = 0;
};

typedef uint16_t precision_t;   // fp16 bit pattern; the host never does arithmetic on it

// colour / camera helpers (internal/utils/common_device.h:106-120, common_device.cu:16-30)
inline float fov_to_focal_length(int resolution, float rad) { return 0.5f * (float)resolution / tanf(0.5f * rad); }
inline float focal_length_to_fov(int resolution, float focal_length) { return 2.f * 180.f / 3.14159265358979323846f * atanf((float)resolution / (focal_length * 2.f)); }

// 3x4 camera-to-world, column-major (12 floats: columns 0..2 rotation, column 3 origin), as Eigen stores it
struct Mat34 { float m[12] = {0}; float& at(int r, int c) { return m[c * 3 + r]; } float at(int r, int c) const { return m[c * 3 + r]; } };

inline Mat34 spec_opengl_to_opencv(const Mat34& ogl, float scale, const float offset[3]) {
	Mat34 o = ogl;
	for (int r = 0; r < 3; ++r) { o.at(r, 1) *= -1.0f; o.at(r, 2) *= -1.0f; o.at(r, 3) = o.at(r, 3) * scale + offset[r]; }
	Mat34 c;
	for (int col = 0; col < 4; ++col) { c.at(0, col) = o.at(1, col); c.at(1, col) = o.at(2, col); c.at(2, col) = o.at(0, col); }
	return c;
}

} // namespace ngp
