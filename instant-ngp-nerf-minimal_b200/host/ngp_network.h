// NGPNetwork + Trainer -- the reference's internal/network/ngp_network.h:148-782 module boundary and the slice of
// tiny-cuda-nn's Trainer / optimizer chain it is driven by (trainer.h:54-347, optimizers/{adam,ema,exponential_decay}.h).
// The hash grid, SH encoding, both MLPs and the glue kernels are ONE fused call into the C ABI per direction.
#pragma once
#include "common.h"
#include "json.h"

namespace ngp {

class NGPNetwork {
public:
	explicit NGPNetwork(Json& config);

	// params: fp16 arena in the reference's order [density MLP | rgb MLP | hash grid] (ngp_network.h:496-541)
	void set_params(precision_t* params, precision_t* inference_params, precision_t* gradients) { m_params = params; m_inference_params = inference_params; m_gradients = gradients; }

	// ngp_network.h:217-284.  input: Coordinate[n]; output: [n][4] fp16.  d_n_valid (optional) = device-side element count.
	void inference_mixed_precision(cudaStream_t stream, const Coordinate* input, uint32_t n, const uint32_t* d_n_valid, precision_t* output, bool use_inference_matrices = true, DeviceBuffer<uint8_t>* workspace = nullptr);   // workspace: the caller's own scratch (concurrent calls on several streams); default: the network's
	// ngp_network.h:286-307.  input: positions with the given float stride; output [n][16] fp16
	void density(cudaStream_t stream, const float* positions, uint32_t stride, uint32_t n, precision_t* output, bool use_inference_matrices = true);
	// tiny-cuda-nn/object.h:126-150 as Testbed calls it (testbed.cu:558-560): d output[dim] / d input written over the inputs, with
	// the inference weights; only dim == 3 (sigma_raw, what RenderMode::Normals needs) is implemented
	void input_gradient(cudaStream_t stream, uint32_t dim, Coordinate* input_and_output, uint32_t n, const uint32_t* d_n_valid = nullptr);
	// ngp_network.h:325-398 / 400-494.  forward() only notes its arguments: the fused kernel recomputes the activations in
	// shared memory during backward() instead of storing them in HBM, and writes `output` there.
	void forward(cudaStream_t stream, const Coordinate* input, uint32_t n, precision_t* output);
	// src_index (optional): row of the last inference_mixed_precision() batch (n_encoded rows) each input came from; the encodings
	// computed there are then gathered instead of recomputed (same values: the table has not changed in between).
	void backward(cudaStream_t stream, const Coordinate* input, uint32_t n, const precision_t* output, const precision_t* dL_doutput,
	              const uint32_t* src_index = nullptr, uint32_t n_encoded = 0);
	// first parameter of grid level `level` in the arena
	size_t level_param_begin(uint32_t level) const { return (size_t)NGP_MLP_PARAMS + (size_t)m_grid.offsets[level] * 2; }

	size_t n_params() const { return (size_t)NGP_MLP_PARAMS + (size_t)m_n_grid_entries * 2; }
	size_t n_matrix_params() const { return NGP_MLP_PARAMS; }
	uint32_t padded_output_width() const { return 4; }            // rgbsigma rows are 4 halfs wide (ngp_network.h:68-83)
	uint32_t padded_density_output_width() const { return 16; }
	std::vector<std::pair<uint32_t, uint32_t>> layer_sizes() const { return {{64, 32}, {16, 64}, {64, 32}, {64, 64}, {16, 64}}; }
	const ngp_grid_desc& grid() const { return m_grid; }
	uint32_t n_grid_entries() const { return m_n_grid_entries; }

private:
	void ensure_workspace(uint32_t n);
	ngp_grid_desc m_grid;
	uint32_t m_n_grid_entries = 0;
	precision_t* m_params = nullptr;
	precision_t* m_inference_params = nullptr;
	precision_t* m_gradients = nullptr;
	DeviceBuffer<uint8_t> m_workspace;          // inference / density scratch; starts with the encodings of the last batch
	DeviceBuffer<uint8_t> m_train_workspace;    // backward() scratch, separate so it can read m_workspace's encodings
	precision_t* m_forward_output = nullptr;
};

// Ema(ExponentialDecay(Adam)) as configured by base.json:50-67
struct OptimizerConfig {
	float learning_rate = 1e-2f, beta1 = 0.9f, beta2 = 0.99f, epsilon = 1e-15f, l2_reg = 1e-6f;
	uint32_t decay_start = 20000, decay_interval = 10000; float decay_base = 0.33f; bool has_decay = false;
	float ema_decay = 0.95f; bool has_ema = false;
	static OptimizerConfig from_json(const Json& config);
};

// parameter range [begin, end) rank `rank` of `world` owns in the sharded data-parallel optimizer: equal shards of a multiple of 8
// parameters (16-byte fp16 vectors, in-place all-gather friendly); the last shard absorbs the remainder
struct ShardRange { size_t begin, end, size; };
inline ShardRange shard_range(size_t n_params, uint32_t rank, uint32_t world) {
	const size_t padded = (n_params + 7) / 8 * 8;
	const size_t size = ((padded + world - 1) / world + 7) / 8 * 8;
	return {std::min((size_t)rank * size, padded), std::min((size_t)(rank + 1) * size, padded), size};
}

class Trainer {
public:
	Trainer(std::shared_ptr<NGPNetwork> network, const OptimizerConfig& opt, uint32_t seed, cudaStream_t stream);
	void initialize_params(cudaStream_t stream);
	// Trainer::optimizer_step (trainer.h:198): ExponentialDecay -> Adam -> EMA in one fused launch
	void optimizer_step(cudaStream_t stream, float loss_scale);
	// data parallel over peer memory: this rank updates only [shard_begin(), shard_end()) and stores the new weights into every rank's arena
	void set_shard(uint32_t rank, uint32_t world);
	size_t shard_size() const { return m_shard_size; }
	size_t shard_begin() const { return shard_range(m_n, (uint32_t)m_shard_rank, (uint32_t)m_shard_world).begin; }
	size_t shard_end() const { return shard_range(m_n, (uint32_t)m_shard_rank, (uint32_t)m_shard_world).end; }
	size_t padded_n() const { return (m_n + 7) / 8 * 8; }
	void optimizer_step_sharded(cudaStream_t stream, float loss_scale, const ngp_peer_set& peers, uint32_t epoch);
	precision_t* ema() const { return m_ema.data(); }
	float* m1() const { return m_m1.data(); }
	float* m2() const { return m_m2.data(); }
	uint32_t* param_steps() const { return m_param_steps.data(); }

	precision_t* params() const { return m_params.data(); }
	precision_t* params_inference() const { return m_opt.has_ema ? m_ema.data() : m_params.data(); }
	precision_t* gradients() const { return m_grads.data(); }
	float* params_full_precision() const { return m_params_fp32.data(); }
	size_t n_params() const { return m_n; }
	uint32_t step() const { return m_step; }

	Json serialize(bool serialize_optimizer);
	void deserialize(const Json& data, cudaStream_t stream);

private:
	std::shared_ptr<NGPNetwork> m_network;
	OptimizerConfig m_opt;
	uint32_t m_seed;
	size_t m_n;
	DeviceBuffer<float> m_params_fp32, m_m1, m_m2;
	DeviceBuffer<precision_t> m_params, m_grads, m_ema;
	DeviceBuffer<uint32_t> m_param_steps;
	DeviceBuffer<uint32_t> m_ticket;
	size_t m_shard_size = 0, m_shard_rank = 0, m_shard_world = 1;
	uint32_t m_step = 0;
	float m_lr_factor = 1.0f;
};

} // namespace ngp
