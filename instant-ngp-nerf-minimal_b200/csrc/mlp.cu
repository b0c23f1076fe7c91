// Fully fused density + colour MLPs on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM) for sm_100a.
// Replaces tiny-cuda-nn's kernel_mlp_fused / kernel_mlp_fused_backward (src/fully_fused_mlp.cu:152,480), the five split-K
// CUTLASS weight-gradient GEMMs and the two FullLayer dgrad GEMMs (:845-897), the SH encoding kernel
// (encodings/spherical_harmonics.h:48) and NGPNetwork's glue kernels (internal/network/ngp_network.h:44-127).
//
// Two kernels (see their headers below): mlp_fwd_kernel (inference / density: activations never leave tensor memory) and
// mlp_train_kernel (forward + dgrad chain through tensor memory, weight gradients resident in TMEM for the CTA's lifetime).
// Both are persistent, one CTA per SM, with several independent 128-sample tile slots in flight; the weights (20 KB) stay in
// shared memory; activations never visit HBM; no split-K workspaces, no stored activations.
#include "common.cuh"
#include "tcgen05.cuh"

namespace ngp {

using namespace tc;

enum { MODE_INFER = 0, MODE_TRAIN = 1, MODE_DENSITY = 2 };

// weight offsets in the arena (halfs) and in shared memory (bytes)
constexpr uint32_t kW0 = 0, kW1 = 2048, kV0 = 3072, kV1 = 5120, kV2 = 9216;

struct Smem {
	// weights, core-matrix tiled, rows = output neuron, F = fan-in (the only shared-memory operand of the chain MMAs)
	static constexpr uint32_t W0 = 0, W1 = 4096, V0 = 6144, V1 = 10240, V2 = 18432, WEND = 20480;
};

__device__ __forceinline__ void stage_weights(const __half* __restrict__ w, uint8_t* smem, uint32_t off, uint32_t rows, uint32_t F) {
	// 16-byte chunks: chunk c of row r -> tiled position
	const uint32_t chunks = F / 8;
	for (uint32_t i = threadIdx.x; i < rows * chunks; i += blockDim.x) {
		const uint32_t r = i / chunks, c = i % chunks;
		*reinterpret_cast<uint4*>(smem + off + tile_off(r, c, F)) = *reinterpret_cast<const uint4*>(w + (size_t)r * F + c * 8);
	}
}

__device__ __forceinline__ uint32_t pack2(float a, float b) { const __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<const uint32_t*>(&h); }

// write 8 consecutive features [8*chunk, 8*chunk+8) of this thread's row
__device__ __forceinline__ void store_chunk(uint8_t* smem, uint32_t buf, uint32_t F, uint32_t row, uint32_t chunk, const float* v, bool relu) {
	float t[8];
	#pragma unroll
	for (int i = 0; i < 8; ++i) t[i] = relu ? fmaxf(v[i], 0.f) : v[i];
	uint4 o; o.x = pack2(t[0], t[1]); o.y = pack2(t[2], t[3]); o.z = pack2(t[4], t[5]); o.w = pack2(t[6], t[7]);
	*reinterpret_cast<uint4*>(smem + buf + tile_off(row, chunk, F)) = o;
}

// degree-4 real spherical harmonics of 2*d-1 (spherical_harmonics.h:63-100)
__device__ __forceinline__ void sh4(float dx, float dy, float dz, float* out) {
	const float x = dx * 2.f - 1.f, y = dy * 2.f - 1.f, z = dz * 2.f - 1.f;
	const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
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
}

// read this thread's row: n16 groups of 16 fp32 columns starting at column col
template <int N16>
__device__ __forceinline__ void read_row(uint32_t tmem, uint32_t col, float* v) {
	const uint32_t lane_base = (threadIdx.x & ~31u) << 16;
	// all groups are requested before the one wait: one TMEM round trip per epilogue instead of N16
	uint32_t r[N16 * 16];
	#pragma unroll
	for (int g = 0; g < N16; ++g) tmem_ld16_nowait(tmem + lane_base + col + g * 16, r + g * 16);
	tmem_ld_wait();
	#pragma unroll
	for (int g = 0; g < N16; ++g) tmem_ld_pin(r + g * 16);
	#pragma unroll
	for (int i = 0; i < N16 * 16; ++i) v[i] = __uint_as_float(r[i]);
}

// =====================================================================================================
// Forward-only kernel (inference on the marched samples, density on the occupancy-refresh positions): activations never leave
// tensor memory.
//
//   * one persistent CTA per SM holds kFwdSlots independent tile slots of 128 samples, one warpgroup (4 warps) per slot; the slots
//     share the 20 KB of weights in shared memory and the 512 TMEM columns (128 per slot) and never meet at a CTA-wide barrier;
//   * the A operand of every layer comes from TENSOR MEMORY: a thread reads its accumulator row (tcgen05.ld), applies ReLU, rounds
//     to fp16, packs two values per 32-bit column and writes the row back with tcgen05.st -- no shared-memory round trip, no
//     generic->async proxy fence, no shared-memory bandwidth spent on activations (the weights are the only shared-memory
//     operand).  Even the encoded position (64 B / sample = 16 packed columns) goes global -> registers -> TMEM;
//   * per slot two 64-column halves ping-pong: layer k reads A from the low columns of one half and accumulates into the other;
//   * per layer a slot synchronises with a 128-thread named barrier (A written), its first thread issues the MMAs and commits to
//     the slot's mbarrier, the warpgroup sleeps on it (try_wait: hardware sleep, ~60 cycles wake-up).  (A dedicated MMA warp that
//     polled the slots' "A written" mbarriers was measured first: mbarrier.test_wait costs ~150 cycles a probe, a layer step took
//     2 100 cycles and the kernel was slower than the one it replaces.)
// =====================================================================================================
constexpr uint32_t kFwdSlots = 4;
constexpr uint32_t kFwdThreads = kFwdSlots * 128;

// round to fp16 first, then one max.f16x2 with zero: 2 instructions per pair instead of 3 (rounding is monotonic, so the result equals
// rounding the fp32 ReLU)
__device__ __forceinline__ uint32_t relu_pack(uint32_t a, uint32_t b) {
	const __half2 h = __hmax2(__floats2half2_rn(__uint_as_float(a), __uint_as_float(b)), __floats2half2_rn(0.f, 0.f));
	return *reinterpret_cast<const uint32_t*>(&h);
}
__device__ __forceinline__ void slot_barrier(uint32_t slot) { asm volatile("bar.sync %0, 128;" :: "r"(slot + 1) : "memory"); }

// accumulator columns [src, src + 64) of this thread's lane -> ReLU -> fp16 pairs -> columns [dst, dst + 32)
__device__ __forceinline__ void relu_row_to_tmem(uint32_t src, uint32_t dst) {
	uint32_t r[64], q[32];
	tmem_ld32_nowait(src, r);
	tmem_ld32_nowait(src + 32, r + 32);
	tmem_ld_wait();
	tmem_ld_pin(r); tmem_ld_pin(r + 16); tmem_ld_pin(r + 32); tmem_ld_pin(r + 48);
	#pragma unroll
	for (int i = 0; i < 32; ++i) q[i] = relu_pack(r[2 * i], r[2 * i + 1]);
	tmem_st16(dst, q);
	tmem_st16(dst + 16, q + 16);
}

template <int MODE>
__global__ void __launch_bounds__(kFwdThreads, 1) mlp_fwd_kernel(
	const __half* __restrict__ enc, const float* __restrict__ coords, uint32_t coord_stride, uint32_t n, const uint32_t* __restrict__ n_valid,
	const __half* __restrict__ weights, __half* __restrict__ out)
{
	extern __shared__ __align__(1024) uint8_t smem[];
	constexpr uint32_t kCtrl = Smem::WEND;            // [0,4) tmem base; from +16: one mbarrier ("accumulator complete") per slot
	const uint32_t sbase = smem_u32(smem);
	const uint32_t warp = threadIdx.x >> 5;
	const uint32_t slot = warp >> 2, row = threadIdx.x & 127;
	const uint32_t acc_ready = sbase + kCtrl + 16 + slot * 8;

	if (row == 0) { mbar_init(acc_ready, 1); fence_mbar_init(); }
	if (warp == 0) tmem_alloc(sbase + kCtrl, 512);
	chain_enter();
	stage_weights(weights + kW0, smem, Smem::W0, 64, 32);
	stage_weights(weights + kW1, smem, Smem::W1, 16, 64);
	if (MODE != MODE_DENSITY) {
		stage_weights(weights + kV0, smem, Smem::V0, 64, 32);
		stage_weights(weights + kV1, smem, Smem::V1, 64, 64);
		stage_weights(weights + kV2, smem, Smem::V2, 16, 64);
	}
	fence_proxy_async();                               // the weights are read by the tensor core (async proxy)
	fence_before_sync();
	__syncthreads();
	fence_after_sync();
	const uint32_t tm = *reinterpret_cast<const uint32_t*>(smem + kCtrl);

	const uint32_t n_limit = n_valid ? min(*n_valid, n) : n;
	const uint32_t n_tiles = (n_limit + 127) / 128;
	const uint32_t tile_stride = gridDim.x * kFwdSlots;
	constexpr uint32_t I_128x64 = instr_desc(128, 64, 0, 0), I_128x16 = instr_desc(128, 16, 0, 0), I_128x8 = instr_desc(128, 8, 0, 0);

	const uint32_t mlo = tm + slot * 128, mhi = mlo + 64;                 // the slot's two halves as the MMA addresses them (lane 0)
	const uint32_t lo = mlo + ((row & ~31u) << 16), hi = lo + 64;         // ... and as this thread's warp does (its 32 lanes)
	uint32_t parity = 0;
	// this warpgroup's TMEM stores are complete and ordered before the MMAs its first thread issues; then everyone sleeps on the commit
	auto run_layer = [&](uint32_t d, uint32_t idesc, uint32_t a, uint32_t w_off, uint32_t F) {
		tmem_st_wait();
		fence_before_sync();
		slot_barrier(slot);
		if (row == 0) {
			fence_after_sync();
			gemm_ts(d, idesc, a, k_major(sbase + w_off, F), F / 16, 0);
			mma_commit(acc_ready);
		}
		mbar_wait(acc_ready, parity);
		parity ^= 1;
		fence_after_sync();
	};

	struct In { uint32_t e[16]; float dx, dy, dz; };
	auto load_tile = [&](uint32_t tile, In& t) {
		const uint32_t s = tile * 128 + row;
		const bool valid = s < n_limit;
		const uint4* src = reinterpret_cast<const uint4*>(enc + (size_t)s * 32);
		#pragma unroll
		for (int c = 0; c < 4; ++c) {
			const uint4 q = valid ? __ldg(src + c) : make_uint4(0, 0, 0, 0);
			t.e[4 * c] = q.x; t.e[4 * c + 1] = q.y; t.e[4 * c + 2] = q.z; t.e[4 * c + 3] = q.w;
		}
		t.dx = t.dy = t.dz = 0.f;
		if (MODE != MODE_DENSITY && valid) { const float* cp = coords + (size_t)s * coord_stride + 4; t.dx = cp[0]; t.dy = cp[1]; t.dz = cp[2]; }
	};
	uint32_t tile = blockIdx.x * kFwdSlots + slot;
	In nxt;
	if (tile < n_tiles) load_tile(tile, nxt);
	for (; tile < n_tiles; tile += tile_stride) {
		const uint32_t s = tile * 128 + row;
		const bool valid = s < n_limit;
		const In cur = nxt;
		// the encoded position is already fp16 pairs in feature order: 16 packed columns, straight into the A operand of layer 0
		tmem_st16(lo, cur.e);
		if (tile + tile_stride < n_tiles) load_tile(tile + tile_stride, nxt);
		run_layer(mhi, I_128x64, mlo, Smem::W0, 32);                       // h = W0 e
		relu_row_to_tmem(hi, hi);
		run_layer(mlo, I_128x16, mhi, Smem::W1, 64);                       // o = W1 h (16 outputs, no activation)
		uint32_t r[16];
		tmem_ld16_nowait(lo, r);
		tmem_ld_wait();
		tmem_ld_pin(r);
		if (MODE == MODE_DENSITY) {
			if (valid) {
				uint4 o0, o1;
				o0.x = pack2(__uint_as_float(r[0]), __uint_as_float(r[1])); o0.y = pack2(__uint_as_float(r[2]), __uint_as_float(r[3]));
				o0.z = pack2(__uint_as_float(r[4]), __uint_as_float(r[5])); o0.w = pack2(__uint_as_float(r[6]), __uint_as_float(r[7]));
				o1.x = pack2(__uint_as_float(r[8]), __uint_as_float(r[9])); o1.y = pack2(__uint_as_float(r[10]), __uint_as_float(r[11]));
				o1.z = pack2(__uint_as_float(r[12]), __uint_as_float(r[13])); o1.w = pack2(__uint_as_float(r[14]), __uint_as_float(r[15]));
				reinterpret_cast<uint4*>(out + (size_t)s * 16)[0] = o0;
				reinterpret_cast<uint4*>(out + (size_t)s * 16)[1] = o1;
			}
			continue;
		}
		const float sigma_raw = __half2float(__float2half_rn(__uint_as_float(r[0])));
		{
			uint32_t x[16];
			#pragma unroll
			for (int i = 0; i < 8; ++i) x[i] = pack2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
			float sh[16];
			if (valid) {
				sh4(cur.dx, cur.dy, cur.dz, sh);
			} else {
				#pragma unroll
				for (int i = 0; i < 16; ++i) sh[i] = 0.f;
			}
			#pragma unroll
			for (int i = 0; i < 8; ++i) x[8 + i] = pack2(sh[2 * i], sh[2 * i + 1]);
			tmem_st16(lo, x);                                              // rgb-network input [density_out(16) | SH(16)]
		}
		run_layer(mhi, I_128x64, mlo, Smem::V0, 32);                       // h1 = V0 x
		relu_row_to_tmem(hi, hi);
		run_layer(mlo, I_128x64, mhi, Smem::V1, 64);                       // h2 = V1 h1
		relu_row_to_tmem(lo, lo);
		run_layer(mhi, I_128x8, mlo, Smem::V2, 64);                        // c = V2 h2 (3 of the 8 computed outputs are used)
		uint32_t c[8];
		tmem_ld8_nowait(hi, c);
		tmem_ld_wait();
		tmem_ld_pin8(c);
		if (valid) {
			uint2 o; o.x = pack2(__uint_as_float(c[0]), __uint_as_float(c[1])); o.y = pack2(__uint_as_float(c[2]), sigma_raw);
			*reinterpret_cast<uint2*>(out + (size_t)s * 4) = o;
		}
	}
	fence_before_sync();
	__syncthreads();
	if (warp == 0) tmem_dealloc(tm, 512);
}

// =====================================================================================================
// Training kernel: forward + backward (dgrad and wgrad) of both networks on one tile of 128 samples per slot.
//
//   * one persistent CTA per SM, kTrainSlots = 3 independent tile slots (one warpgroup each) + one weight-gradient warp;
//   * the CHAIN (5 forward layers, 5 dgrad layers: each needs the previous one's result) runs like the forward-only kernel:
//     the A operand of every chain MMA is written to TENSOR MEMORY by the epilogue (tcgen05.st) and read from there, so the
//     critical path of a layer is  accumulator -> registers -> TMEM -> next MMA  with one 128-thread named barrier;
//   * the SAME packed registers are then stored to shared memory -- off the critical path, while the chain MMA already runs -- as
//     the operands of the weight-gradient GEMMs (dW = dY^T X, K = 128 samples, both operands MN-major views of sample-major
//     buffers).  A dedicated thread issues all weight-gradient MMAs of the CTA (one issuer: the accumulators, resident in TMEM
//     for the CTA's lifetime, are shared by the three slots) as the slots hand their operand pairs over through mbarriers;
//   * per-slot shared memory is 68 KB: the five activations a tile needs until its backward pass (E, H, X, H1, H2: 64 KB) and
//     dL/dc (4 KB); every other gradient overwrites an activation whose weight-gradient GEMM has completed (dh2 and dh take
//     H2's place, dh1 takes H1's, dL/d(density_out) takes dL/dc's) -- which is what lets THREE tiles be in flight per SM
//     instead of two (20 + 3 x 68 = 224 KB);
//   * TMEM: per slot 96 columns (A operand 32 | accumulator 64), weight-gradient accumulators 160 columns: 448 of 512.
// =====================================================================================================
constexpr uint32_t kTrainSlots = 3;
constexpr uint32_t kTrainThreads = kTrainSlots * 128 + 32;

struct TrainSmem {
	static constexpr uint32_t E = 0, H = 8192, X = 24576, H1 = 32768, H2 = 49152, G16 = 65536, SLOT = 69632;
	static constexpr uint32_t SLOT0 = Smem::WEND, CTRL = SLOT0 + kTrainSlots * SLOT, END = CTRL + 256;
};
constexpr uint32_t kTrainWg = kTrainSlots * 96;       // first weight-gradient accumulator column
constexpr uint32_t kWgW0 = kTrainWg, kWgW1 = kTrainWg + 32, kWgV0 = kTrainWg + 48, kWgV1 = kTrainWg + 80, kWgV2 = kTrainWg + 144;

// this thread's row of `n_chunks` 16-byte chunks (8 fp16 features each, already packed) into a core-matrix tiled buffer
template <int N_CHUNKS>
__device__ __forceinline__ void store_row_packed(uint8_t* smem, uint32_t buf, uint32_t F, uint32_t row, const uint32_t* q) {
	#pragma unroll
	for (int c = 0; c < N_CHUNKS; ++c) *reinterpret_cast<uint4*>(smem + buf + tile_off(row, c, F)) = make_uint4(q[4 * c], q[4 * c + 1], q[4 * c + 2], q[4 * c + 3]);
}

// accumulator columns [src, src + 64) -> fp16 pairs q[32]; RELU: max with zero; else: zero where the forward activation (this
// thread's row of the 64-wide buffer `act`) was not positive
template <bool RELU>
__device__ __forceinline__ void row64_to_packed(uint32_t src, uint32_t* q, const uint8_t* smem = nullptr, uint32_t act = 0, uint32_t row = 0) {
	uint32_t r[64];
	tmem_ld32_nowait(src, r);
	tmem_ld32_nowait(src + 32, r + 32);
	tmem_ld_wait();
	tmem_ld_pin(r); tmem_ld_pin(r + 16); tmem_ld_pin(r + 32); tmem_ld_pin(r + 48);
	if (RELU) {
		#pragma unroll
		for (int i = 0; i < 32; ++i) q[i] = relu_pack(r[2 * i], r[2 * i + 1]);
	} else {
		const __half2 zero = __floats2half2_rn(0.f, 0.f);
		#pragma unroll
		for (int c = 0; c < 8; ++c) {
			const uint4 a = *reinterpret_cast<const uint4*>(smem + act + tile_off(row, c, 64));
			const uint32_t w[4] = {a.x, a.y, a.z, a.w};
			#pragma unroll
			for (int i = 0; i < 4; ++i) {
				const __half2 g = __floats2half2_rn(__uint_as_float(r[8 * c + 2 * i]), __uint_as_float(r[8 * c + 2 * i + 1]));
				const __half2 m = __hmul2(g, __hgt2(*reinterpret_cast<const __half2*>(&w[i]), zero));      // gradient x (activation > 0)
				q[4 * c + i] = *reinterpret_cast<const uint32_t*>(&m);
			}
		}
	}
}

__global__ void __launch_bounds__(kTrainThreads, 1) mlp_train_kernel(
	const __half* __restrict__ enc, const float* __restrict__ coords, uint32_t coord_stride, uint32_t n, const uint32_t* __restrict__ n_valid,
	const __half* __restrict__ weights, const __half* __restrict__ dloss, __half* __restrict__ out, __half* __restrict__ denc, float* __restrict__ wgrad,
	float* __restrict__ wgrad_partials, const uint32_t* __restrict__ enc_rows, uint32_t n_enc_rows)
{
	extern __shared__ __align__(1024) uint8_t smem[];
	using TS = TrainSmem;
	const uint32_t sbase = smem_u32(smem);
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const bool wg_warp = warp == kTrainSlots * 4;
	// control block: [0,4) tmem base; barriers (8 B each) from +16: acc_ready[slot], wg_ready[slot], wg_done[slot]
	auto bar_acc = [&](uint32_t s) { return sbase + TS::CTRL + 16 + s * 8; };
	auto bar_wg_ready = [&](uint32_t s) { return sbase + TS::CTRL + 16 + (kTrainSlots + s) * 8; };
	auto bar_wg_done = [&](uint32_t s) { return sbase + TS::CTRL + 16 + (2 * kTrainSlots + s) * 8; };

	if (threadIdx.x == kTrainSlots * 128) {
		for (uint32_t s = 0; s < kTrainSlots; ++s) { mbar_init(bar_acc(s), 1); mbar_init(bar_wg_ready(s), 4); mbar_init(bar_wg_done(s), 1); }
		fence_mbar_init();
	}
	if (wg_warp) tmem_alloc(sbase + TS::CTRL, 512);
	chain_enter();
	stage_weights(weights + kW0, smem, Smem::W0, 64, 32);
	stage_weights(weights + kW1, smem, Smem::W1, 16, 64);
	stage_weights(weights + kV0, smem, Smem::V0, 64, 32);
	stage_weights(weights + kV1, smem, Smem::V1, 64, 64);
	stage_weights(weights + kV2, smem, Smem::V2, 16, 64);
	fence_proxy_async();
	fence_before_sync();
	__syncthreads();
	fence_after_sync();
	const uint32_t tm = *reinterpret_cast<const uint32_t*>(smem + TS::CTRL);

	const uint32_t n_limit = n_valid ? min(*n_valid, n) : n;
	const uint32_t n_tiles = (n_limit + 127) / 128;
	const uint32_t tile_stride = gridDim.x * kTrainSlots;

	if (wg_warp) {
		// ---------------------------------------------------------------------------------------------------------------
		// weight-gradient issuer: per tile round five jobs (V2, V1, V0, W1, W0) for each slot in turn; every job waits for the slot's
		// operand pair ("wg_ready": the gradient has been stored and fenced by all four warps), issues dW += dY^T X (M = 64,
		// K = 128 samples) and commits to the slot's "wg_done", which the slot waits on before it reuses either buffer
		// ---------------------------------------------------------------------------------------------------------------
		if (lane == 0) {
			constexpr uint32_t I_w64 = instr_desc(64, 64, 1, 1), I_w32 = instr_desc(64, 32, 1, 1), I_w16 = instr_desc(64, 16, 1, 1);
			// FIXED order (tile round, job, slot): every accumulator sums its tiles in the same order whatever the timing, so the weight
			// gradients are deterministic; the slots run the same sequence side by side, so waiting in order costs little
			uint32_t parity = 0, started = 0;                                      // started: bit k = accumulator k holds a sum
			for (uint32_t base = blockIdx.x * kTrainSlots; base < n_tiles; base += tile_stride) {
				#pragma unroll 1
				for (uint32_t job = 0; job < 5; ++job) {
					#pragma unroll
					for (uint32_t s = 0; s < kTrainSlots; ++s) {
						if (base + s >= n_tiles) continue;
						mbar_wait(bar_wg_ready(s), parity);
						fence_after_sync();
						const uint32_t b = sbase + TS::SLOT0 + s * TS::SLOT;
						const uint32_t acc = (started >> job) & 1u;
						switch (job) {
							case 0: gemm(tm + kWgV2, I_w16, mn_major(b + TS::H2, 64), mn_major(b + TS::G16, 16), 8, acc); break;      // dV2^T[64][16] += H2^T dc
							case 1: gemm(tm + kWgV1, I_w64, mn_major(b + TS::H2, 64), mn_major(b + TS::H1, 64), 8, acc); break;       // dV1[64][64]  += dh2^T H1   (dh2 in H2's place)
							case 2: gemm(tm + kWgV0, I_w32, mn_major(b + TS::H1, 64), mn_major(b + TS::X, 32), 8, acc); break;        // dV0[64][32]  += dh1^T x    (dh1 in H1's place)
							case 3: gemm(tm + kWgW1, I_w16, mn_major(b + TS::H, 64), mn_major(b + TS::G16, 16), 8, acc); break;       // dW1^T[64][16] += H^T do    (do in dc's place)
							default: gemm(tm + kWgW0, I_w32, mn_major(b + TS::H2, 64), mn_major(b + TS::E, 32), 8, acc); break;       // dW0[64][32]  += dh^T e     (dh in H2's place)
						}
						mma_commit(bar_wg_done(s));
						started |= 1u << job;
					}
					parity ^= 1;                                                   // every slot's hand-over barrier completes one phase per job
				}
			}
		}
		__syncwarp();
	} else {
		const uint32_t slot = warp >> 2, row = threadIdx.x & 127;
		const uint32_t acc_ready = bar_acc(slot), wg_ready = bar_wg_ready(slot), wg_done = bar_wg_done(slot);
		const uint32_t sb = TS::SLOT0 + slot * TS::SLOT;                      // this slot's buffers (byte offset in smem[])
		const uint32_t mA = tm + slot * 96, mD = mA + 32;                      // A operand / accumulator columns as the MMA addresses them
		const uint32_t tA = mA + ((row & ~31u) << 16), tD = tA + 32;           // ... and as this thread's warp does (its 32 lanes)
		uint32_t parity = 0, wg_parity = 0;
		constexpr uint32_t I_128x64 = instr_desc(128, 64, 0, 0), I_128x16 = instr_desc(128, 16, 0, 0);
		constexpr uint32_t I_d64 = instr_desc(128, 64, 0, 1), I_d32 = instr_desc(128, 32, 0, 1), I_d16 = instr_desc(128, 16, 0, 1);   // dgrad: B = W as MN-major

		// chain step, part 1: the A operand (this warpgroup's TMEM stores) is complete -> the slot's first thread issues the layer
		auto issue = [&](uint32_t idesc, Operand b, uint32_t nk) {
			tmem_st_wait();
			fence_before_sync();
			slot_barrier(slot);
			if (row == 0) {
				fence_after_sync();
				gemm_ts(mD, idesc, mA, b, nk, 0);
				mma_commit(acc_ready);
			}
		};
		// chain step, part 2: sleep until the accumulator is complete
		auto acquire = [&]() { mbar_wait(acc_ready, parity); parity ^= 1; fence_after_sync(); };
		// the shared-memory stores of this thread are visible to the tensor core; the fourth warp's arrival releases the weight-gradient job
		auto hand_over = [&]() { fence_proxy_async(); __syncwarp(); if (lane == 0) mbar_arrive(wg_ready); };
		auto wait_wgrad = [&]() { mbar_wait(wg_done, wg_parity); wg_parity ^= 1; };

		struct In { uint32_t e[16]; float dx, dy, dz; uint2 dl; };
		auto load_tile = [&](uint32_t tile, In& t) {
			const uint32_t s = tile * 128 + row;
			const bool valid = s < n_limit;
			const uint32_t enc_row = (enc_rows && valid) ? min(__ldg(enc_rows + s), n_enc_rows - 1u) : s;       // reuse of the inference pass's encodings
			const uint4* src = reinterpret_cast<const uint4*>(enc + (size_t)enc_row * 32);
			#pragma unroll
			for (int c = 0; c < 4; ++c) {
				const uint4 v = valid ? __ldg(src + c) : make_uint4(0, 0, 0, 0);
				t.e[4 * c] = v.x; t.e[4 * c + 1] = v.y; t.e[4 * c + 2] = v.z; t.e[4 * c + 3] = v.w;
			}
			t.dx = t.dy = t.dz = 0.f;
			if (valid) { const float* cp = coords + (size_t)s * coord_stride + 4; t.dx = cp[0]; t.dy = cp[1]; t.dz = cp[2]; }
			// dloss == nullptr: "input gradient" mode (tiny-cuda-nn/object.h:126-150) -- the loss gradient is one-hot on sigma_raw with the
			// reference's backprop scale 128 (fp16 0x5800), zero on r, g, b
			t.dl = make_uint2(0u, 0u);
			if (valid) t.dl = dloss ? __ldg(reinterpret_cast<const uint2*>(dloss + (size_t)s * 4)) : make_uint2(0u, 0x58000000u);
		};
		uint32_t tile = blockIdx.x * kTrainSlots + slot;
		In nxt;
		if (tile < n_tiles) load_tile(tile, nxt);
		for (; tile < n_tiles; tile += tile_stride) {
			const uint32_t s = tile * 128 + row;
			const bool valid = s < n_limit;
			const In cur = nxt;
			uint32_t q[32];
			// ================= forward =================
			tmem_st16(tA, cur.e);
			issue(I_128x64, k_major(sbase + Smem::W0, 32), 2);                                       // h = W0 e
			store_row_packed<4>(smem, sb + TS::E, 32, row, cur.e);
			if (tile + tile_stride < n_tiles) load_tile(tile + tile_stride, nxt);
			acquire();
			row64_to_packed<true>(tD, q);
			tmem_st16(tA, q); tmem_st16(tA + 16, q + 16);
			issue(I_128x16, k_major(sbase + Smem::W1, 64), 4);                                       // o = W1 h
			store_row_packed<8>(smem, sb + TS::H, 64, row, q);
			acquire();
			float sigma_raw;
			{
				uint32_t r[16];
				tmem_ld16_nowait(tD, r);
				tmem_ld_wait();
				tmem_ld_pin(r);
				sigma_raw = __half2float(__float2half_rn(__uint_as_float(r[0])));
				#pragma unroll
				for (int i = 0; i < 8; ++i) q[i] = pack2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
				float sh[16];
				if (valid) {
					sh4(cur.dx, cur.dy, cur.dz, sh);
				} else {
					#pragma unroll
					for (int i = 0; i < 16; ++i) sh[i] = 0.f;
				}
				#pragma unroll
				for (int i = 0; i < 8; ++i) q[8 + i] = pack2(sh[2 * i], sh[2 * i + 1]);
			}
			tmem_st16(tA, q);
			issue(I_128x64, k_major(sbase + Smem::V0, 32), 2);                                       // h1 = V0 [o | sh]
			store_row_packed<4>(smem, sb + TS::X, 32, row, q);
			acquire();
			row64_to_packed<true>(tD, q);
			tmem_st16(tA, q); tmem_st16(tA + 16, q + 16);
			issue(I_128x64, k_major(sbase + Smem::V1, 64), 4);                                       // h2 = V1 h1
			store_row_packed<8>(smem, sb + TS::H1, 64, row, q);
			acquire();
			row64_to_packed<true>(tD, q);
			tmem_st16(tA, q); tmem_st16(tA + 16, q + 16);
			issue(I_128x16, k_major(sbase + Smem::V2, 64), 4);                                       // c = V2 h2
			store_row_packed<8>(smem, sb + TS::H2, 64, row, q);
			acquire();
			float dsigma = 0.f;
			{
				uint32_t c[8];
				tmem_ld8_nowait(tD, c);
				tmem_ld_wait();
				tmem_ld_pin8(c);
				if (valid && out) {
					uint2 o; o.x = pack2(__uint_as_float(c[0]), __uint_as_float(c[1])); o.y = pack2(__uint_as_float(c[2]), sigma_raw);
					*reinterpret_cast<uint2*>(out + (size_t)s * 4) = o;
				}
				// ================= backward =================
				// dL/dc: r, g, b from the loss; the 13 padding outputs get no gradient
				const __half2 hi = *reinterpret_cast<const __half2*>(&cur.dl.y);
				dsigma = __half2float(__high2half(hi));
				const __half2 b0 = __halves2half2(__low2half(hi), __float2half_rn(0.f));
				#pragma unroll
				for (int i = 0; i < 8; ++i) q[i] = 0u;
				q[0] = cur.dl.x; q[1] = *reinterpret_cast<const uint32_t*>(&b0);
			}
			tmem_st8(tA, q);
			issue(I_d64, mn_major(sbase + Smem::V2, 64), 1);                                         // dh2 = dc V2
			store_row_packed<2>(smem, sb + TS::G16, 16, row, q);
			hand_over();                                                                             // job V2: (H2, dc)
			acquire();
			row64_to_packed<false>(tD, q, smem, sb + TS::H2, row);                                   // x relu'(h2)
			tmem_st16(tA, q); tmem_st16(tA + 16, q + 16);
			issue(I_d64, mn_major(sbase + Smem::V1, 64), 4);                                         // dh1 = dh2 V1
			wait_wgrad();                                                                            // job V2 has read H2 ...
			store_row_packed<8>(smem, sb + TS::H2, 64, row, q);                                    // ... whose place dh2 takes
			hand_over();                                                                             // job V1: (dh2, H1)
			acquire();
			row64_to_packed<false>(tD, q, smem, sb + TS::H1, row);                                   // x relu'(h1)
			tmem_st16(tA, q); tmem_st16(tA + 16, q + 16);
			issue(I_d16, mn_major(sbase + Smem::V0, 32), 4);                                         // dx[0..16) = dh1 V0[:, 0..16): the SH half has no parameters
			wait_wgrad();                                                                            // job V1 has read H1 (and dh2)
			store_row_packed<8>(smem, sb + TS::H1, 64, row, q);                                    // dh1 takes H1's place
			hand_over();                                                                             // job V0: (dh1, x)
			acquire();
			{
				uint32_t r[16];
				tmem_ld16_nowait(tD, r);
				tmem_ld_wait();
				tmem_ld_pin(r);
				float v0 = __half2float(__float2half_rn(__uint_as_float(r[0]))) + dsigma;              // ngp_network.h:119-126
				q[0] = pack2(v0, __uint_as_float(r[1]));
				#pragma unroll
				for (int i = 1; i < 8; ++i) q[i] = pack2(__uint_as_float(r[2 * i]), __uint_as_float(r[2 * i + 1]));
			}
			tmem_st8(tA, q);
			issue(I_d64, mn_major(sbase + Smem::W1, 64), 1);                                         // dh = do W1
			store_row_packed<2>(smem, sb + TS::G16, 16, row, q);                                   // do takes dc's place (job V2 is long complete)
			wait_wgrad();                                                                            // job V0 has been taken: the hand-over barrier never runs a phase ahead of its reader
			hand_over();                                                                             // job W1: (H, do)
			acquire();
			row64_to_packed<false>(tD, q, smem, sb + TS::H, row);                                    // x relu'(h)
			tmem_st16(tA, q); tmem_st16(tA + 16, q + 16);
			issue(I_d32, mn_major(sbase + Smem::W0, 32), 4);                                         // de = dh W0
			store_row_packed<8>(smem, sb + TS::H2, 64, row, q);                                    // dh takes dh2's place (job V1 is complete)
			wait_wgrad();                                                                            // job W1 has been taken
			hand_over();                                                                             // job W0: (dh, e)
			acquire();
			{
				uint32_t r[32];
				tmem_ld32_nowait(tD, r);
				tmem_ld_wait();
				tmem_ld_pin(r); tmem_ld_pin(r + 16);
				if (valid && denc) {
					uint4* dst = reinterpret_cast<uint4*>(denc + (size_t)s * 32);
					#pragma unroll
					for (int c = 0; c < 4; ++c) {
						uint4 o;
						o.x = pack2(__uint_as_float(r[8 * c]), __uint_as_float(r[8 * c + 1])); o.y = pack2(__uint_as_float(r[8 * c + 2]), __uint_as_float(r[8 * c + 3]));
						o.z = pack2(__uint_as_float(r[8 * c + 4]), __uint_as_float(r[8 * c + 5])); o.w = pack2(__uint_as_float(r[8 * c + 6]), __uint_as_float(r[8 * c + 7]));
						dst[c] = o;
					}
				}
			}
			// the next tile overwrites E (and the flush reads the accumulators): job W0 must be complete
			wait_wgrad();
		}
	}

	// ---- flush the weight-gradient accumulators (M = 64: row m sits in TMEM lane 32*(m/16) + m%16) ------
	// wgrad_partials != nullptr: this CTA owns one [NGP_MLP_PARAMS] fp32 slot and stores plainly (a small second kernel sums
	// the slots in a fixed order); otherwise fp32 atomics into wgrad (the stand-alone ngp_mlp_train entry point).
	fence_before_sync();
	__syncthreads();
	fence_after_sync();
	if (blockIdx.x * kTrainSlots >= n_tiles) {
		// no tile (a device-side count shrank the batch): this CTA's slot still takes part in the fixed-order sum
		if (wgrad_partials) for (uint32_t i = threadIdx.x; i < NGP_MLP_PARAMS; i += blockDim.x) wgrad_partials[(size_t)blockIdx.x * NGP_MLP_PARAMS + i] = 0.f;
	} else if (warp < 4) {
		const uint32_t m = warp * 16 + lane;
		float v[64];
		if (wgrad_partials) {
			float* slot_out = wgrad_partials + (size_t)blockIdx.x * NGP_MLP_PARAMS;
			read_row<2>(tm, kWgW0, v);      if (lane < 16) for (int k = 0; k < 32; k += 4) *reinterpret_cast<float4*>(slot_out + kW0 + m * 32 + k) = make_float4(v[k], v[k + 1], v[k + 2], v[k + 3]);
			read_row<1>(tm, kWgW1, v);      if (lane < 16) for (int o = 0; o < 16; ++o) slot_out[kW1 + o * 64 + m] = v[o];
			read_row<2>(tm, kWgV0, v);      if (lane < 16) for (int k = 0; k < 32; k += 4) *reinterpret_cast<float4*>(slot_out + kV0 + m * 32 + k) = make_float4(v[k], v[k + 1], v[k + 2], v[k + 3]);
			read_row<4>(tm, kWgV1, v);      if (lane < 16) for (int k = 0; k < 64; k += 4) *reinterpret_cast<float4*>(slot_out + kV1 + m * 64 + k) = make_float4(v[k], v[k + 1], v[k + 2], v[k + 3]);
			read_row<1>(tm, kWgV2, v);      if (lane < 16) for (int o = 0; o < 16; ++o) slot_out[kV2 + o * 64 + m] = v[o];
		} else {
			read_row<2>(tm, kWgW0, v);      if (lane < 16) for (int k = 0; k < 32; ++k) atomicAdd(wgrad + kW0 + m * 32 + k, v[k]);
			read_row<1>(tm, kWgW1, v);      if (lane < 16) for (int o = 0; o < 16; ++o) atomicAdd(wgrad + kW1 + o * 64 + m, v[o]);
			read_row<2>(tm, kWgV0, v);      if (lane < 16) for (int k = 0; k < 32; ++k) atomicAdd(wgrad + kV0 + m * 32 + k, v[k]);
			read_row<4>(tm, kWgV1, v);      if (lane < 16) for (int k = 0; k < 64; ++k) atomicAdd(wgrad + kV1 + m * 64 + k, v[k]);
			read_row<1>(tm, kWgV2, v);      if (lane < 16) for (int o = 0; o < 16; ++o) atomicAdd(wgrad + kV2 + o * 64 + m, v[o]);
		}
	}
	fence_before_sync();
	__syncthreads();
	if (wg_warp) tmem_dealloc(tm, 512);
}

} // namespace ngp

// =====================================================================================================
// host side
// =====================================================================================================
using namespace ngp;

namespace ngp {
int hash_encode_fwd_counted(cudaStream_t stream, const float* d_xyz, uint32_t stride_floats, uint32_t n, const uint32_t* d_n_valid,
                            const void* d_table, const ngp_grid_desc* desc, void* d_out, uint32_t out_stride);

// sums the per-CTA weight-gradient slots and rounds to the fp16 gradient arena.  Deterministic: thread (i, j) of a CTA adds the
// slots j, j + 8, j + 16, ... of parameter i in that order, then the 8 partial sums are added in the order of j -- the same
// tree whatever the grid.  (One thread per parameter walking all 296 slots took 15 us: 80 CTAs, one load in flight each.)
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ partials, uint32_t n_slots, __half* __restrict__ w16, uint32_t n) {
	chain_enter();
	__shared__ float part[8][32];
	const uint32_t li = threadIdx.x & 31, j = threadIdx.x >> 5, i = blockIdx.x * 32 + li;
	float s = 0.f;
	if (i < n) {
		for (uint32_t c = j; c < n_slots; c += 8) s += partials[(size_t)c * n + i];
	}
	part[j][li] = s;
	__syncthreads();
	if (j == 0 && i < n) {
		float t = part[0][li];
		#pragma unroll
		for (int k = 1; k < 8; ++k) t += part[k][li];
		w16[i] = __float2half_rn(t);
	}
}

template <int MODE>
static int launch_mlp(cudaStream_t stream, const void* enc, const float* coords, uint32_t coord_stride, uint32_t n, const uint32_t* n_valid,
                      const void* weights, const void* dloss, void* out, void* denc, float* wgrad, float* wgrad_partials = nullptr, uint32_t* grid_out = nullptr,
                      const uint32_t* enc_rows = nullptr, uint32_t n_enc_rows = 0) {
	if (MODE != MODE_TRAIN) {
		// forward only: the warp-specialised kernel with activations in tensor memory, one persistent CTA per SM
		constexpr uint32_t fwd_smem = Smem::WEND + 256;
		static bool fwd_configured = false;
		if (!fwd_configured) {
			if (cudaFuncSetAttribute(mlp_fwd_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, fwd_smem) != cudaSuccess) {
				set_error("mlp: cannot reserve %u bytes of shared memory: %s", fwd_smem, cudaGetErrorString(cudaGetLastError())); return -2;
			}
			fwd_configured = true;
		}
		const uint32_t groups = div_up(div_up(n, 128), kFwdSlots);
		const uint32_t fwd_grid = groups < kNumSMs ? groups : kNumSMs;
		NGP_LAUNCH("mlp_fwd_kernel", stream, launch_chained(mlp_fwd_kernel<MODE>, fwd_grid, kFwdThreads, fwd_smem, stream, (const __half*)enc, coords, coord_stride, n, n_valid,
			(const __half*)weights, (__half*)out));
		if (grid_out) *grid_out = fwd_grid;
		return check_launch("mlp_fwd_kernel");
	}
	// training: forward + backward, three tile slots per persistent CTA
	static bool train_configured = false;
	if (!train_configured) {
		if (cudaFuncSetAttribute(mlp_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, TrainSmem::END) != cudaSuccess) {
			set_error("mlp: cannot reserve %u bytes of shared memory: %s", TrainSmem::END, cudaGetErrorString(cudaGetLastError())); return -2;
		}
		train_configured = true;
	}
	const uint32_t groups = div_up(div_up(n, 128), kTrainSlots);
	const uint32_t grid = groups < kNumSMs ? groups : kNumSMs;
	NGP_LAUNCH("mlp_kernel", stream, launch_chained(mlp_train_kernel, grid, kTrainThreads, TrainSmem::END, stream, (const __half*)enc, coords, coord_stride, n, n_valid, (const __half*)weights,
		(const __half*)dloss, (__half*)out, (__half*)denc, wgrad, wgrad_partials, enc_rows, n_enc_rows));
	if (grid_out) *grid_out = grid;
	return check_launch("mlp_kernel");
}
} // namespace ngp

namespace ngp {
int hash_input_gradient(cudaStream_t stream, float* d_coords, uint32_t n, const uint32_t* d_n_valid, const void* d_table, const ngp_grid_desc* desc, const void* d_denc, float inv_scale);
}

// NGPNetwork::input_gradient(stream, 3, coords, coords) (tiny-cuda-nn/object.h:126-150 through ngp_network.h:325-494): d sigma_raw /
// d input, written OVER the coordinates like the reference does (testbed.cu:558-560 passes the same matrix as input and output):
// position <- gradient, warped dt <- dt / 128 (the reference rescales the whole matrix, including the row its backward pass never
// wrote), direction <- 0.  Uses the weights it is given (the reference uses the inference / EMA weights).
extern "C" int ngp_network_input_gradient(void* stream, float* d_coords, uint32_t n, const uint32_t* d_n_valid, const void* d_params,
                                          const ngp_grid_desc* desc, void* d_workspace, size_t workspace_bytes) {
	NGP_REQUIRE(d_coords && d_params && desc && d_workspace, "null pointer");
	NGP_REQUIRE(desc->n_levels == 16, "the fused network expects 16 levels x 2 features = 32 encoded dims");
	NGP_REQUIRE(workspace_bytes >= ngp_network_workspace_bytes(n), "workspace too small");
	if (n == 0) return 0;
	cudaStream_t s = (cudaStream_t)stream;
	const __half* params = (const __half*)d_params;
	const size_t n_pad = ((size_t)n + 127) / 128 * 128;
	__half* enc = (__half*)d_workspace;
	__half* denc = enc + n_pad * 32;
	float* wgrad_slots = (float*)(denc + n_pad * 32);       // the weight gradients of this pass are not wanted; the slots absorb them
	int rc = hash_encode_fwd_counted(s, d_coords, 7, n, d_n_valid, params + NGP_MLP_PARAMS, desc, enc, 32);
	if (rc) return rc;
	rc = launch_mlp<MODE_TRAIN>(s, enc, d_coords, 7, n, d_n_valid, params, nullptr, nullptr, denc, nullptr, wgrad_slots);
	if (rc) return rc;
	return hash_input_gradient(s, d_coords, n, d_n_valid, params + NGP_MLP_PARAMS, desc, denc, 1.0f / 128.0f);
}

extern "C" int ngp_mlp_inference(void* stream, const void* d_enc, const float* d_coords, uint32_t n, const void* d_mlp_params, void* d_rgbsigma) {
	NGP_REQUIRE(d_enc && d_coords && d_mlp_params && d_rgbsigma, "null pointer");
	if (n == 0) return 0;
	return launch_mlp<MODE_INFER>((cudaStream_t)stream, d_enc, d_coords, 7, n, nullptr, d_mlp_params, nullptr, d_rgbsigma, nullptr, nullptr);
}

extern "C" int ngp_mlp_train(void* stream, const void* d_enc, const float* d_coords, uint32_t n, const void* d_mlp_params, const void* d_dloss,
                             void* d_rgbsigma, void* d_denc, float* d_wgrad_f32) {
	NGP_REQUIRE(d_enc && d_coords && d_mlp_params && d_dloss && d_wgrad_f32, "null pointer");
	if (n == 0) return 0;
	return launch_mlp<MODE_TRAIN>((cudaStream_t)stream, d_enc, d_coords, 7, n, nullptr, d_mlp_params, d_dloss, d_rgbsigma, d_denc, d_wgrad_f32);
}

// workspace: [enc n*32 halfs][denc n*32 halfs][per-CTA weight-gradient slots, kMaxTrainCtas x 10240 floats]
constexpr uint32_t kMaxTrainCtas = kNumSMs * 2;
extern "C" size_t ngp_network_workspace_bytes(uint32_t n) {
	const size_t n_pad = ((size_t)n + 127) / 128 * 128;
	return n_pad * 32 * 2 * 2 + (size_t)kMaxTrainCtas * NGP_MLP_PARAMS * sizeof(float) + 256;
}

extern "C" int ngp_network_inference(void* stream, const float* d_coords, uint32_t n, const uint32_t* d_n_valid, const void* d_params,
                                     const ngp_grid_desc* desc, void* d_rgbsigma, void* d_workspace, size_t workspace_bytes) {
	NGP_REQUIRE(d_coords && d_params && desc && d_rgbsigma && d_workspace, "null pointer");
	NGP_REQUIRE(desc->n_levels == 16, "the fused network expects 16 levels x 2 features = 32 encoded dims");
	NGP_REQUIRE(workspace_bytes >= ngp_network_workspace_bytes(n), "workspace too small");
	if (n == 0) return 0;
	cudaStream_t s = (cudaStream_t)stream;
	const __half* params = (const __half*)d_params;
	int rc = hash_encode_fwd_counted(s, d_coords, 7, n, d_n_valid, params + NGP_MLP_PARAMS, desc, d_workspace, 32);
	if (rc) return rc;
	return launch_mlp<MODE_INFER>(s, d_workspace, d_coords, 7, n, d_n_valid, params, nullptr, d_rgbsigma, nullptr, nullptr);
}

extern "C" int ngp_network_density(void* stream, const float* d_positions, uint32_t stride_floats, uint32_t n, const void* d_params,
                                   const ngp_grid_desc* desc, void* d_out, void* d_workspace, size_t workspace_bytes) {
	NGP_REQUIRE(d_positions && d_params && desc && d_out && d_workspace, "null pointer");
	NGP_REQUIRE(desc->n_levels == 16, "the fused network expects 16 levels x 2 features = 32 encoded dims");
	NGP_REQUIRE(workspace_bytes >= ngp_network_workspace_bytes(n), "workspace too small");
	if (n == 0) return 0;
	cudaStream_t s = (cudaStream_t)stream;
	const __half* params = (const __half*)d_params;
	int rc = hash_encode_fwd_counted(s, d_positions, stride_floats, n, nullptr, params + NGP_MLP_PARAMS, desc, d_workspace, 32);
	if (rc) return rc;
	return launch_mlp<MODE_DENSITY>(s, d_workspace, nullptr, 0, n, nullptr, params, nullptr, d_out, nullptr, nullptr);
}

static int network_train_impl(void* stream, const float* d_coords, uint32_t n, const void* d_params, const ngp_grid_desc* desc, const void* d_dloss,
                              void* d_rgbsigma, void* d_grads, const void* d_encoded, uint32_t n_encoded, const uint32_t* d_src_index, void* d_workspace, size_t workspace_bytes,
                              int zero_first = 1) {
	NGP_REQUIRE(d_coords && d_params && desc && d_dloss && d_grads && d_workspace, "null pointer");
	NGP_REQUIRE(desc->n_levels == 16, "the fused network expects 16 levels x 2 features = 32 encoded dims");
	NGP_REQUIRE(workspace_bytes >= ngp_network_workspace_bytes(n), "workspace too small");
	if (n == 0) return 0;
	cudaStream_t s = (cudaStream_t)stream;
	const __half* params = (const __half*)d_params;
	const size_t n_pad = ((size_t)n + 127) / 128 * 128;
	__half* enc = (__half*)d_workspace;
	__half* denc = enc + n_pad * 32;
	float* wgrad_slots = (float*)(denc + n_pad * 32);
	int rc = 0;
	if (!d_encoded) rc = hash_encode_fwd_counted(s, d_coords, 7, n, nullptr, params + NGP_MLP_PARAMS, desc, enc, 32);
	if (rc) return rc;
	uint32_t n_slots = 0;
	rc = launch_mlp<MODE_TRAIN>(s, d_encoded ? d_encoded : enc, d_coords, 7, n, nullptr, params, d_dloss, d_rgbsigma, denc, nullptr, wgrad_slots, &n_slots, d_src_index, n_encoded);
	if (rc) return rc;
	NGP_LAUNCH("wgrad_reduce_kernel", s, launch_chained(wgrad_reduce_kernel, div_up(NGP_MLP_PARAMS, 32), 256, 0, s, wgrad_slots, n_slots, (__half*)d_grads, NGP_MLP_PARAMS));
	__half* grid_grads = (__half*)d_grads + NGP_MLP_PARAMS;
	return ngp_hash_encode_bwd(stream, d_coords, 7, n, denc, 32, desc, grid_grads, zero_first);
}

extern "C" int ngp_network_train(void* stream, const float* d_coords, uint32_t n, const void* d_params, const ngp_grid_desc* desc,
                                 const void* d_dloss, void* d_rgbsigma, void* d_grads, void* d_workspace, size_t workspace_bytes) {
	return network_train_impl(stream, d_coords, n, d_params, desc, d_dloss, d_rgbsigma, d_grads, nullptr, 0, nullptr, d_workspace, workspace_bytes);
}

extern "C" int ngp_network_train_reuse(void* stream, const float* d_coords, uint32_t n, const void* d_params, const ngp_grid_desc* desc,
                                       const void* d_dloss, void* d_rgbsigma, void* d_grads, const void* d_encoded, uint32_t n_encoded, const uint32_t* d_src_index,
                                       void* d_workspace, size_t workspace_bytes) {
	NGP_REQUIRE(d_encoded && d_src_index, "null pointer");
	NGP_REQUIRE(n_encoded > 0, "n_encoded must be positive");
	NGP_REQUIRE(d_encoded != d_workspace, "the training workspace must not alias the encodings it reads");
	return network_train_impl(stream, d_coords, n, d_params, desc, d_dloss, d_rgbsigma, d_grads, d_encoded, n_encoded, d_src_index, d_workspace, workspace_bytes);
}
