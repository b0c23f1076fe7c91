#include "ngp_network.h"
#include <algorithm>
#include <cctype>

namespace ngp {

static std::string lower(std::string s) { for (auto& c : s) c = (char)tolower(c); return s; }

NGPNetwork::NGPNetwork(Json& config) {
	Json& enc = config["encoding"];
	// auto_determine_grid_parameters (ngp_network.h:597-641)
	const std::string otype = lower(enc.value("otype", "OneBlob"));
	if (otype.find("grid") == std::string::npos) throw std::runtime_error("NGPNetwork: only the HashGrid position encoding is implemented (got '" + otype + "')");
	const uint32_t n_features_per_level = enc.value("n_features_per_level", 2u);
	uint32_t n_levels = enc.value("n_levels", 16u);
	if (enc.contains("n_features") && enc["n_features"].number() > 0) n_levels = (uint32_t)enc["n_features"].number() / n_features_per_level;
	const uint32_t log2_hashmap_size = enc.value("log2_hashmap_size", 15u);
	uint32_t base_resolution = enc.value("base_resolution", 0u);
	if (!base_resolution) { base_resolution = 1u << (log2_hashmap_size / 3); enc["base_resolution"] = Json(base_resolution); }
	const float desired_resolution = enc.value("desired_resolution", 2048.0f);
	float per_level_scale = enc.value("per_level_scale", 0.0f);
	if (per_level_scale <= 0.0f && n_levels > 1) { per_level_scale = ngp_auto_per_level_scale(desired_resolution, base_resolution, n_levels); enc["per_level_scale"] = Json(per_level_scale); }
	if (n_features_per_level != 2 || n_levels != 16)
		throw std::runtime_error("NGPNetwork: the fused sm_100a network is built for 16 levels x 2 features (32 encoded dims)");
	if (lower(enc.value("interpolation", "Linear")) != "linear") throw std::runtime_error("NGPNetwork: only linear grid interpolation is implemented");
	m_n_grid_entries = ngp_grid_desc_init(&m_grid, n_levels, log2_hashmap_size, base_resolution, per_level_scale);
	if (!m_n_grid_entries) throw std::runtime_error(ngp_last_error());
	// the per-level scales the reference's kernels evaluate on the device (grid.h:157), not glibc's: an ulp apart on some levels
	DeviceBuffer<float> scratch(NGP_MAX_LEVELS);
	api_check(ngp_grid_desc_device_scales(&m_grid, base_resolution, per_level_scale, nullptr, scratch.data()));

	auto check_mlp = [](const Json& c, int hidden, const char* name) {
		if (lower(c.value("otype", "FullyFusedMLP")) != "fullyfusedmlp" || c.value("n_neurons", 64) != 64 || c.value("n_hidden_layers", hidden) != hidden ||
		    lower(c.value("activation", "ReLU")) != "relu" || lower(c.value("output_activation", "None")) != "none")
			throw std::runtime_error(std::string("NGPNetwork: ") + name + " must be a 64-wide ReLU FullyFusedMLP as in base.json (no other otype is implemented)");
	};
	check_mlp(config["network"], 1, "network");
	check_mlp(config["rgb_network"], 2, "rgb_network");
	const Json& dir = config["dir_encoding"];
	if (lower(dir.value("otype", "SphericalHarmonics")) != "sphericalharmonics" || dir.value("degree", 4) != 4)
		throw std::runtime_error("NGPNetwork: dir_encoding must be SphericalHarmonics of degree 4");
}

void NGPNetwork::ensure_workspace(uint32_t n) { m_workspace.enlarge(ngp_network_workspace_bytes(n)); }

void NGPNetwork::inference_mixed_precision(cudaStream_t stream, const Coordinate* input, uint32_t n, const uint32_t* d_n_valid, precision_t* output, bool use_inference_matrices,
                                           DeviceBuffer<uint8_t>* workspace) {
	DeviceBuffer<uint8_t>& ws = workspace ? *workspace : m_workspace;
	ws.enlarge(ngp_network_workspace_bytes(n));
	api_check(ngp_network_inference(stream, (const float*)input, n, d_n_valid, use_inference_matrices ? m_inference_params : m_params, &m_grid, output,
	                                ws.data(), ws.get_bytes()));
}

void NGPNetwork::density(cudaStream_t stream, const float* positions, uint32_t stride, uint32_t n, precision_t* output, bool use_inference_matrices) {
	ensure_workspace(n);
	api_check(ngp_network_density(stream, positions, stride, n, use_inference_matrices ? m_inference_params : m_params, &m_grid, output,
	                              m_workspace.data(), m_workspace.get_bytes()));
}

void NGPNetwork::input_gradient(cudaStream_t stream, uint32_t dim, Coordinate* input_and_output, uint32_t n, const uint32_t* d_n_valid) {
	if (dim >= padded_output_width()) throw std::runtime_error("Invalid dimension to compute the input gradient for.");
	if (dim != 3) throw std::runtime_error("input_gradient: only the density output (dim 3) is implemented");
	m_train_workspace.enlarge(ngp_network_workspace_bytes(n));
	api_check(ngp_network_input_gradient(stream, (float*)input_and_output, n, d_n_valid, m_inference_params, &m_grid, m_train_workspace.data(), m_train_workspace.get_bytes()));
}

void NGPNetwork::forward(cudaStream_t, const Coordinate*, uint32_t, precision_t* output) { m_forward_output = output; }

void NGPNetwork::backward(cudaStream_t stream, const Coordinate* input, uint32_t n, const precision_t* output, const precision_t* dL_doutput,
                          const uint32_t* src_index, uint32_t n_encoded) {
	precision_t* out = m_forward_output == output ? m_forward_output : nullptr;
	m_train_workspace.enlarge(ngp_network_workspace_bytes(n));
	if (src_index) {
		// the batch is a subset of the samples inference_mixed_precision just encoded into m_workspace: gather instead of re-encoding
		api_check(ngp_network_train_reuse(stream, (const float*)input, n, m_params, &m_grid, dL_doutput, out, m_gradients, m_workspace.data(), n_encoded, src_index,
		                                  m_train_workspace.data(), m_train_workspace.get_bytes()));
	} else {
		api_check(ngp_network_train(stream, (const float*)input, n, m_params, &m_grid, dL_doutput, out, m_gradients, m_train_workspace.data(), m_train_workspace.get_bytes()));
	}
	m_forward_output = nullptr;
}

OptimizerConfig OptimizerConfig::from_json(const Json& config) {
	OptimizerConfig o;
	const Json* c = &config;
	for (int depth = 0; depth < 4; ++depth) {
		const std::string t = lower(c->value("otype", "Adam"));
		if (t == "ema") { o.has_ema = true; o.ema_decay = c->value("decay", 0.99f); c = &c->at("nested"); }
		else if (t == "exponentialdecay") {
			o.has_decay = true; o.decay_base = c->value("decay_base", 0.33f); o.decay_start = c->value("decay_start", 10000u);
			o.decay_interval = c->value("decay_interval", 10000u); c = &c->at("nested");
		} else if (t == "adam") {
			o.learning_rate = c->value("learning_rate", 1e-3f); o.beta1 = c->value("beta1", 0.9f); o.beta2 = c->value("beta2", 0.999f);
			o.epsilon = c->value("epsilon", 1e-8f); o.l2_reg = c->value("l2_reg", 1e-8f);
			return o;
		} else throw std::runtime_error("optimizer otype '" + t + "' is not implemented (Ema / ExponentialDecay / Adam are)");
	}
	throw std::runtime_error("optimizer nesting too deep");
}

Trainer::Trainer(std::shared_ptr<NGPNetwork> network, const OptimizerConfig& opt, uint32_t seed, cudaStream_t stream)
	: m_network(network), m_opt(opt), m_seed(seed), m_n(network->n_params()) {
	// padding: the 16-byte accesses of the fused optimizer never leave the allocation, and world x shard_size (<= n + 8 * world)
	// fits for the in-place all-gathers of the data-parallel mode
	const size_t n_alloc = m_n + 8 + 8 * NGP_MAX_PEERS;
	m_ticket.resize(1); m_ticket.memset(0, stream);
	m_shard_size = (m_n + 7) / 8 * 8;
	m_params_fp32.resize(n_alloc); m_m1.resize(n_alloc); m_m2.resize(n_alloc);
	m_params.resize(n_alloc); m_grads.resize(n_alloc); m_ema.resize(n_alloc); m_param_steps.resize(n_alloc);
	m_network->set_params(m_params.data(), params_inference(), m_grads.data());
	initialize_params(stream);
}

void Trainer::initialize_params(cudaStream_t stream) {
	// trainer.h:54-62: pcg32 seeded from std::seed_seq{seed}
	std::seed_seq seq{m_seed};
	std::vector<uint32_t> seeds(2);
	seq.generate(seeds.begin(), seeds.end());
	Pcg32 rng{(uint64_t)seeds.front()};
	m_params_fp32.memset(0, stream); m_params.memset(0, stream);
	uint64_t state = rng.state;
	api_check(ngp_init_params(stream, &state, rng.inc, (uint64_t)m_network->n_grid_entries() * 2, m_params_fp32.data(), m_params.data()));
	m_m1.memset(0, stream); m_m2.memset(0, stream); m_param_steps.memset(0, stream); m_grads.memset(0, stream);
	// EMA shadow weights start equal to the weights (ema.h: first step debiases from zero; inference before any step uses params)
	NGP_CUDA_CHECK(cudaMemcpyAsync(m_ema.data(), m_params.data(), m_params.get_bytes(), cudaMemcpyDeviceToDevice, stream));
	m_step = 0; m_lr_factor = 1.0f;
	NGP_CUDA_CHECK(cudaStreamSynchronize(stream));
}

void Trainer::optimizer_step(cudaStream_t stream, float loss_scale) {
	// exponential_decay.h:62-73, evaluated with the pre-increment step count
	if (m_opt.has_decay && m_step >= m_opt.decay_start && (m_step - m_opt.decay_start) % m_opt.decay_interval == 0) m_lr_factor *= m_opt.decay_base;
	ngp_optimizer_cfg c{m_opt.learning_rate, m_opt.beta1, m_opt.beta2, m_opt.epsilon, m_opt.l2_reg, loss_scale, m_opt.has_ema ? m_opt.ema_decay : 0.0f};
	api_check(ngp_optimizer_step(stream, &c, m_n, m_network->n_matrix_params(), m_step, m_opt.learning_rate * m_lr_factor, m_params_fp32.data(), m_params.data(),
	                             m_grads.data(), m_m1.data(), m_m2.data(), m_param_steps.data(), m_ema.data()));
	++m_step;
}

void Trainer::set_shard(uint32_t rank, uint32_t world) {
	if (world < 1 || world > NGP_MAX_PEERS || rank >= world) throw std::runtime_error("Trainer::set_shard: bad rank / world");
	m_shard_rank = rank; m_shard_world = world;
	m_shard_size = shard_range(m_n, rank, world).size;
}

void Trainer::optimizer_step_sharded(cudaStream_t stream, float loss_scale, const ngp_peer_set& peers, uint32_t epoch) {
	if (m_opt.has_decay && m_step >= m_opt.decay_start && (m_step - m_opt.decay_start) % m_opt.decay_interval == 0) m_lr_factor *= m_opt.decay_base;
	ngp_optimizer_cfg c{m_opt.learning_rate, m_opt.beta1, m_opt.beta2, m_opt.epsilon, m_opt.l2_reg, loss_scale, m_opt.has_ema ? m_opt.ema_decay : 0.0f};
	api_check(ngp_optimizer_step_sharded(stream, &c, shard_begin(), shard_end(), m_network->n_matrix_params(), m_step, m_opt.learning_rate * m_lr_factor,
	                                     m_params_fp32.data(), m_m1.data(), m_m2.data(), m_param_steps.data(), m_ema.data(), &peers, epoch, m_ticket.data()));
	++m_step;
}

// trainer.h:294-317 + the optimizer serialisers, key for key: Ema {nested, weights_ema_binary} (ema.h:174-185), ExponentialDecay
// {nested, learning_rate, learning_rate_factor} (exponential_decay.h:118-129), Adam {current_step, base_learning_rate,
// first_moments_binary, second_moments_binary, param_steps_binary} (adam.h:255-277).  A file written here loads in the reference
// and the reverse.  The reference drops the fp32 master and the fp16 training weights (after a load it resumes from the widened
// EMA weights); they are added under two keys the reference never looks up, so a snapshot written AND read here resumes exactly.
Json Trainer::serialize(bool serialize_optimizer) {
	Json data = Json::object();
	data["n_params"] = Json((uint64_t)m_n);
	std::vector<precision_t> h(m_n);
	(m_opt.has_ema ? m_ema : m_params).copy_to_host(h.data(), m_n);            // m_params_inference (trainer.h:299)
	data["params_binary"] = Json::binary(h.data(), m_n * sizeof(precision_t));
	if (serialize_optimizer) {
		Json node = Json::object();                                             // Adam
		node["current_step"] = Json(m_step);
		node["base_learning_rate"] = Json(m_opt.learning_rate);
		std::vector<float> f(m_n);
		m_m1.copy_to_host(f.data(), m_n); node["first_moments_binary"] = Json::binary(f.data(), m_n * 4);
		m_m2.copy_to_host(f.data(), m_n); node["second_moments_binary"] = Json::binary(f.data(), m_n * 4);
		std::vector<uint32_t> u(m_n);
		m_param_steps.copy_to_host(u.data(), m_n); node["param_steps_binary"] = Json::binary(u.data(), m_n * 4);
		m_params_fp32.copy_to_host(f.data(), m_n); node["params_full_precision_binary"] = Json::binary(f.data(), m_n * 4);   // extra (see above)
		std::vector<precision_t> w(m_n);
		m_params.copy_to_host(w.data(), m_n); node["params_training_binary"] = Json::binary(w.data(), m_n * 2);               // extra
		if (m_opt.has_decay) {
			Json decay = Json::object();
			decay["nested"] = node;
			decay["learning_rate"] = Json(m_opt.learning_rate);
			decay["learning_rate_factor"] = Json(m_lr_factor);
			node = decay;
		}
		if (m_opt.has_ema) {
			Json ema = Json::object();
			ema["nested"] = node;
			ema["weights_ema_binary"] = Json::binary(h.data(), m_n * sizeof(precision_t));
			node = ema;
		}
		data["optimizer"] = node;
	}
	return data;
}

static float half_bits_to_float(uint16_t h) {
	const uint32_t s = (h >> 15) & 1, e = (h >> 10) & 0x1f, m = h & 0x3ff;
	const float v = e == 0 ? std::ldexp((float)m, -24) : e == 31 ? (m ? NAN : INFINITY) : std::ldexp((float)(m | 0x400), (int)e - 25);
	return s ? -v : v;
}

void Trainer::deserialize(const Json& data, cudaStream_t stream) {
	const Json& pb = data.at("params_binary");
	// Trainer::set_params (trainer.h:267-288): the buffer size is what is checked, "n_params" itself is never read back
	if (pb.type != Json::Binary || pb.s.size() != m_n * sizeof(precision_t)) throw std::runtime_error("Can't set params because buffer has the wrong size.");
	NGP_CUDA_CHECK(cudaMemcpyAsync(m_ema.data(), pb.s.data(), pb.s.size(), cudaMemcpyHostToDevice, stream));
	NGP_CUDA_CHECK(cudaMemcpyAsync(m_params.data(), pb.s.data(), pb.s.size(), cudaMemcpyHostToDevice, stream));
	// fp32 master = widened fp16 (set_params)
	std::vector<float> f(m_n);
	const uint16_t* h = (const uint16_t*)pb.s.data();
	for (size_t i = 0; i < m_n; ++i) f[i] = half_bits_to_float(h[i]);
	NGP_CUDA_CHECK(cudaMemcpyAsync(m_params_fp32.data(), f.data(), m_n * 4, cudaMemcpyHostToDevice, stream));
	NGP_CUDA_CHECK(cudaStreamSynchronize(stream));
	if (!data.contains("optimizer")) return;
	auto load = [&](const Json& node, const char* key, void* dst, size_t bytes) {
		const Json& b = node.at(key);
		if (b.type != Json::Binary || b.s.size() != bytes) throw std::runtime_error(std::string("snapshot: ") + key + " has the wrong size");
		NGP_CUDA_CHECK(cudaMemcpy(dst, b.s.data(), bytes, cudaMemcpyHostToDevice));
	};
	const Json* node = &data.at("optimizer");
	if (m_opt.has_ema) {
		load(*node, "weights_ema_binary", m_ema.data(), m_n * 2);                 // ema.h:180 requires it
		node = &node->at("nested");
	}
	if (m_opt.has_decay) {
		(void)node->at("learning_rate");                                          // exponential_decay.h:125 requires it; the value is the configured base rate
		m_lr_factor = node->contains("learning_rate_factor") ? (float)node->at("learning_rate_factor").number() : 1.0f;
		node = &node->at("nested");
	}
	m_step = (uint32_t)node->at("current_step").number();
	(void)node->at("base_learning_rate");
	load(*node, "first_moments_binary", m_m1.data(), m_n * 4);
	load(*node, "second_moments_binary", m_m2.data(), m_n * 4);
	if (node->contains("param_steps_binary")) load(*node, "param_steps_binary", m_param_steps.data(), m_n * 4);
	else m_param_steps.memset(0, stream);                                         // adam.h:268-271
	// our own extras: exact resume.  A reference-written file has neither; training then continues from the widened EMA weights,
	// which is what the reference itself does after load_snapshot.
	if (node->contains("params_full_precision_binary")) load(*node, "params_full_precision_binary", m_params_fp32.data(), m_n * 4);
	if (node->contains("params_training_binary")) load(*node, "params_training_binary", m_params.data(), m_n * 2);
	NGP_CUDA_CHECK(cudaStreamSynchronize(stream));
}

} // namespace ngp
