// Testbed host orchestration.  Follows the schedule of the reference's internal/testbed.cu (file:line cited per method);
// all device work is stream-ordered calls into the C ABI, and the only host synchronisations are the ones whose results
// Python reads (counters + loss at the end of train(), the live-ray count while rendering).
#include "testbed.h"
#include <nccl.h>
#include <algorithm>
#include <chrono>
#include <fstream>
#include <sstream>
#include <sys/stat.h>

namespace ngp {

static std::string file_extension(const std::string& p) {
	const size_t k = p.find_last_of('.');
	std::string e = k == std::string::npos ? "" : p.substr(k + 1);
	for (auto& c : e) c = (char)tolower(c);
	return e;
}
static bool path_exists(const std::string& p) { struct stat st; return !p.empty() && stat(p.c_str(), &st) == 0; }

Testbed::Testbed() {
	// testbed.cu:17-37: a CUDA device with tensor cores must be present; here that means sm_100
	cudaDeviceProp props;
	int dev = 0;
	cudaError_t error = cudaGetDevice(&dev);
	if (error == cudaSuccess) error = cudaGetDeviceProperties(&props, dev);
	if (error != cudaSuccess) throw std::runtime_error(std::string("cudaGetDeviceProperties() returned an error: ") + cudaGetErrorString(error));
	if (props.major != 10) throw std::runtime_error("ngp_b200 kernels are built for sm_100a (B200) only; there is no fallback path.");
	// the training stream outranks the look-ahead march stream: when both have blocks pending, the step's own kernels go first
	int prio_lo = 0, prio_hi = 0;
	NGP_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
	NGP_CUDA_CHECK(cudaStreamCreateWithPriority(&m_stream, cudaStreamNonBlocking, prio_hi));
	for (auto& e : m_ev) NGP_CUDA_CHECK(cudaEventCreate(&e));
	NGP_CUDA_CHECK(cudaStreamCreateWithPriority(&m_march_stream, cudaStreamNonBlocking, prio_lo));
	for (int b = 0; b < 2; ++b) {
		NGP_CUDA_CHECK(cudaStreamCreateWithPriority(&m_band_stream[b], cudaStreamNonBlocking, prio_hi));
		NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_band_done[b], cudaEventDisableTiming));
	}
	NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_frame_cleared, cudaEventDisableTiming));
	NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_march_go, cudaEventDisableTiming));
	for (int k = 0; k < 2; ++k) {
		NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_march_ready[k], cudaEventDisableTiming));
		NGP_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_slot_free[k], cudaEventDisableTiming));
	}
	NGP_CUDA_CHECK(cudaMallocHost((void**)&m_host_counter, 64 * sizeof(uint32_t)));
	m_training_buffer.loss_sum.resize(1024);
	m_training_buffer.loss_sum.memset(0);
	m_rendering_buffer.n_hit_counter.memset(0);
	m_rendering_buffer.n_alive_counter.memset(0);
}

void Testbed::close_peer_handles() {
	for (void* p : m_peer_opened) cudaIpcCloseMemHandle(p);
	m_peer_opened.clear();
	m_peers = ngp_peer_set{};
	m_peer_enabled = false;
}

Testbed::~Testbed() {
	close_peer_handles();
	for (auto& e : m_ev) if (e) cudaEventDestroy(e);
	for (cudaEvent_t e : {m_ev_march_go, m_ev_march_ready[0], m_ev_march_ready[1], m_ev_slot_free[0], m_ev_slot_free[1]}) if (e) cudaEventDestroy(e);
	if (m_march_stream) cudaStreamDestroy(m_march_stream);
	for (int b = 0; b < 2; ++b) { if (m_band_stream[b]) cudaStreamDestroy(m_band_stream[b]); if (m_ev_band_done[b]) cudaEventDestroy(m_ev_band_done[b]); }
	if (m_ev_frame_cleared) cudaEventDestroy(m_ev_frame_cleared);
	if (m_host_counter) cudaFreeHost(m_host_counter);
	if (m_stream) cudaStreamDestroy(m_stream);
}

// testbed.cu:44-93
void Testbed::load_model_config(const std::string& config_path) {
	if (!path_exists(config_path)) throw std::runtime_error("The model config must exist.");
	std::ifstream f(config_path, std::ios::in | std::ios::binary);
	std::stringstream ss; ss << f.rdbuf();
	const std::string ext = file_extension(config_path);
	if (ext == "json") load_model_config_json(Json::parse(ss.str()));
	else if (ext == "msgpack") load_model_config_json(Json::from_msgpack(ss.str()));
	else throw std::runtime_error("The model config must be a .json or .msgpack file.");
}

void Testbed::load_model_config_json(const Json& config) {
	if (m_dp_world > 1) {
		// The Trainer below owns new gradient / parameter arenas: this rank's peer set would point at freed memory and the handles the
		// other ranks opened are stale.  Data-parallel training has to be set up again (collectively) before the next train().
		if (m_stream) cudaStreamSynchronize(m_stream);
		close_peer_handles();
		m_dp_stale = true;
	}
	m_model_config = config;
	m_sampler = std::make_shared<OccupancySampler>(m_model_config["samlper"]);   // (sic) the key the reference reads, testbed.cu:67
	m_network = std::make_shared<NGPNetwork>(m_model_config["network"]);
	m_render = std::make_shared<RayMarcher>(m_model_config["render"]);
	m_loss_type = string_to_loss_type(m_model_config["loss"].str());
	m_trainer = std::make_shared<Trainer>(m_network, OptimizerConfig::from_json(m_model_config["optimizer"]), m_seed, m_stream);
	m_train_envmap = m_model_config.contains("envmap") ? m_model_config["envmap"].value("train", false) : false;
	if (m_train_envmap) {
		// testbed.cu:83-90.  All buffers start at zero (trainable_buffer.h:45-55), so an untrained map changes nothing.
		const Json& e = m_model_config["envmap"];
		const int h = e.value("height", 256), w = e.value("width", 256);
		if (h <= 0 || w <= 0) throw std::runtime_error("envmap: height and width must be positive");
		if (m_dp_world > 1) throw std::runtime_error("envmap training is single-GPU (its gradients are not exchanged)");
		m_envmap.resolution[0] = h; m_envmap.resolution[1] = w;
		m_envmap.loss_type = string_to_loss_type(e.contains("loss") ? e["loss"].str() : std::string("RelativeL2"));
		m_envmap.opt = OptimizerConfig::from_json(e["optimizer"]);
		m_envmap.n_params = (size_t)h * w * 4;
		const size_t n_alloc = m_envmap.n_params + 16;
		m_envmap.params_fp32.resize(n_alloc); m_envmap.m1.resize(n_alloc); m_envmap.m2.resize(n_alloc); m_envmap.param_steps.resize(n_alloc);
		m_envmap.params.resize(n_alloc); m_envmap.gradients.resize(n_alloc); m_envmap.ema.resize(n_alloc);
		m_envmap.params_fp32.memset(0, m_stream); m_envmap.m1.memset(0, m_stream); m_envmap.m2.memset(0, m_stream); m_envmap.param_steps.memset(0, m_stream);
		m_envmap.params.memset(0, m_stream); m_envmap.gradients.memset(0, m_stream); m_envmap.ema.memset(0, m_stream);
		m_envmap.step = 0; m_envmap.lr_factor = 1.0f;
		NGP_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	}
	m_training_buffer.i_step = 0;
}

void Testbed::envmap_optimizer_step(cudaStream_t stream, float loss_scale) {
	Envmap& e = m_envmap;
	if (e.opt.has_decay && e.step >= e.opt.decay_start && (e.step - e.opt.decay_start) % e.opt.decay_interval == 0) e.lr_factor *= e.opt.decay_base;
	ngp_optimizer_cfg c{e.opt.learning_rate, e.opt.beta1, e.opt.beta2, e.opt.epsilon, e.opt.l2_reg, loss_scale, e.opt.has_ema ? e.opt.ema_decay : 0.0f};
	// a TrainableBuffer has no matrix parameters (layer_sizes() is empty): no weight decay, texels without gradient are skipped
	api_check(ngp_optimizer_step(stream, &c, e.n_params, 0, e.step, e.opt.learning_rate * e.lr_factor, e.params_fp32.data(), e.params.data(), e.gradients.data(),
	                             e.m1.data(), e.m2.data(), e.param_steps.data(), e.ema.data()));
	++e.step;
}

// testbed.cu:95-125
void Testbed::load_training_data(const std::string& data_path, float scale, const float offset[3]) {
	if (!m_sampler) throw std::runtime_error("load_model_config must be called before load_training_data");
	if (!path_exists(data_path)) throw std::runtime_error("The data path must exist.");
	if (file_extension(data_path) != "json") throw std::runtime_error("The data path must be a .json file.");
	m_training_buffer.dataset.load_from_json(data_path, scale, offset);
	m_scene_scale = scale;
	for (int k = 0; k < 3; ++k) m_scene_offset[k] = offset[k];
	const NeRFSynthetic& d = m_training_buffer.dataset;
	m_sampler->init(d.n_images, d.image_resolution, d.focal_length, d.principal_point, d.xforms, m_rng, m_stream);
}

void Testbed::load_training_data_from_memory(uint32_t n_images, int width, int height, float camera_angle_x, const Mat34* xforms_ogl, const uint8_t* rgba8,
                                             float scale, const float offset[3]) {
	if (!m_sampler) throw std::runtime_error("load_model_config must be called before load_training_data_from_memory");
	m_training_buffer.dataset.load_from_memory(n_images, width, height, camera_angle_x, xforms_ogl, rgba8, scale, offset);
	m_scene_scale = scale;
	for (int k = 0; k < 3; ++k) m_scene_offset[k] = offset[k];
	const NeRFSynthetic& d = m_training_buffer.dataset;
	m_sampler->init(d.n_images, d.image_resolution, d.focal_length, d.principal_point, d.xforms, m_rng, m_stream);
}

void Testbed::load_procedural_training_data(uint32_t n_images, int width, int height, uint32_t seed, float scale, const float offset[3]) {
	if (!m_sampler) throw std::runtime_error("load_model_config must be called before load_procedural_training_data");
	m_training_buffer.dataset.load_procedural(n_images, width, height, 0.6911112f, seed, scale, offset);
	m_scene_scale = scale;
	for (int k = 0; k < 3; ++k) m_scene_offset[k] = offset[k];
	const NeRFSynthetic& d = m_training_buffer.dataset;
	m_sampler->init(d.n_images, d.image_resolution, d.focal_length, d.principal_point, d.xforms, m_rng, m_stream);
}

// testbed.cu:127-141
void Testbed::save_snapshot(const std::string& snapshot_path, bool serialize_optimizer) {
	if (serialize_optimizer && m_peer_enabled && m_sharded_state_dirty)
		throw std::runtime_error("save_snapshot: the optimizer state is sharded over the ranks; call gather_optimizer_state() on every rank first");
	NGP_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	m_model_config["snapshot"] = m_trainer->serialize(serialize_optimizer);
	std::vector<float> grid(m_sampler->n_total_elements);
	m_sampler->density_grid.copy_to_host(grid.data(), grid.size());
	m_model_config["snapshot"]["density_grid"] = Json::binary(grid.data(), grid.size() * sizeof(float));
	m_model_config["snapshot"]["scene_scale"] = Json(m_scene_scale);
	Json off = Json::array();
	for (int k = 0; k < 3; ++k) off.push_back(Json(m_scene_offset[k]));
	m_model_config["snapshot"]["scene_offset"] = off;
	std::string out;
	m_model_config.to_msgpack(out);
	std::ofstream f(snapshot_path, std::ios::out | std::ios::binary);
	if (!f) throw std::runtime_error("cannot write snapshot: " + snapshot_path);
	f.write(out.data(), (std::streamsize)out.size());
}

// testbed.cu:143-163
void Testbed::load_snapshot(const std::string& snapshot_path) {
	load_model_config(snapshot_path);
	if (!m_model_config.contains("snapshot")) throw std::runtime_error("ERROR in loading snapshot.");
	const Json& snap = m_model_config["snapshot"];
	m_scene_scale = (float)snap.at("scene_scale").number();
	for (int k = 0; k < 3; ++k) m_scene_offset[k] = (float)snap.at("scene_offset")[k].number();
	const Json& grid = snap.at("density_grid");
	if (grid.type != Json::Binary || grid.s.size() != m_sampler->n_total_elements * sizeof(float)) throw std::runtime_error("snapshot: density_grid has the wrong size");
	NGP_CUDA_CHECK(cudaMemcpy(m_sampler->density_grid.data(), grid.s.data(), grid.s.size(), cudaMemcpyHostToDevice));
	m_sampler->update_grid_bitfield(m_stream);
	m_trainer->deserialize(snap, m_stream);
	NGP_CUDA_CHECK(cudaStreamSynchronize(m_stream));
}

void Testbed::set_data_parallel(int rank, int world_size, void* nccl_comm, bool peer_memory) {
	if (world_size < 1 || rank < 0 || rank >= world_size) throw std::runtime_error("set_data_parallel: bad rank / world size");
	if (world_size > 1 && !nccl_comm) throw std::runtime_error("set_data_parallel: a communicator is required for world_size > 1");
	if (!m_trainer) throw std::runtime_error("set_data_parallel: load_model_config first");
	if (m_train_envmap && world_size > 1) throw std::runtime_error("envmap training is single-GPU (its gradients are not exchanged)");
	m_dp_rank = rank; m_dp_world = world_size; m_nccl_comm = nccl_comm;
	close_peer_handles();
	m_dp_stale = false;
	m_sharded_state_dirty = false;
	m_render->set_regulariser_scale(1.f / (float)world_size);
	if (!peer_memory || world_size == 1) return;
	if (world_size > NGP_MAX_PEERS) throw std::runtime_error("set_data_parallel: peer memory supports at most 8 ranks");
	NGP_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	m_peer_flags.resize(2 * NGP_MAX_PEERS);
	m_peer_flags.memset(0);
	// exchange CUDA IPC handles of (gradient arena, parameter arena, flag block) through the communicator
	struct Handles { cudaIpcMemHandle_t grads, params, flags; };
	Handles mine;
	NGP_CUDA_CHECK(cudaIpcGetMemHandle(&mine.grads, m_trainer->gradients()));
	NGP_CUDA_CHECK(cudaIpcGetMemHandle(&mine.params, m_trainer->params()));
	NGP_CUDA_CHECK(cudaIpcGetMemHandle(&mine.flags, m_peer_flags.data()));
	DeviceBuffer<uint8_t> d_all((size_t)world_size * sizeof(Handles));
	NGP_CUDA_CHECK(cudaMemcpy(d_all.data() + (size_t)rank * sizeof(Handles), &mine, sizeof(Handles), cudaMemcpyHostToDevice));
	ncclResult_t r = ncclAllGather(d_all.data() + (size_t)rank * sizeof(Handles), d_all.data(), sizeof(Handles), ncclUint8, (ncclComm_t)nccl_comm, m_stream);
	if (r != ncclSuccess) throw std::runtime_error(std::string("ncclAllGather failed: ") + ncclGetErrorString(r));
	NGP_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	std::vector<Handles> all(world_size);
	d_all.copy_to_host((uint8_t*)all.data(), d_all.get_bytes());
	m_peers = ngp_peer_set{};
	m_peers.world = (uint32_t)world_size; m_peers.rank = (uint32_t)rank;
	for (int p = 0; p < world_size; ++p) {
		if (p == rank) { m_peers.grads[p] = m_trainer->gradients(); m_peers.params[p] = m_trainer->params(); m_peers.flags[p] = m_peer_flags.data(); continue; }
		void *g = nullptr, *w = nullptr, *f = nullptr;
		NGP_CUDA_CHECK(cudaIpcOpenMemHandle(&g, all[p].grads, cudaIpcMemLazyEnablePeerAccess));
		NGP_CUDA_CHECK(cudaIpcOpenMemHandle(&w, all[p].params, cudaIpcMemLazyEnablePeerAccess));
		NGP_CUDA_CHECK(cudaIpcOpenMemHandle(&f, all[p].flags, cudaIpcMemLazyEnablePeerAccess));
		m_peers.grads[p] = g; m_peers.params[p] = w; m_peers.flags[p] = (uint32_t*)f;
		m_peer_opened.push_back(g); m_peer_opened.push_back(w); m_peer_opened.push_back(f);
	}
	m_trainer->set_shard((uint32_t)rank, (uint32_t)world_size);
	m_peer_epoch = 0;
	m_peer_enabled = true;
	// nobody may publish into a flag block that its owner has not zeroed yet
	DeviceBuffer<float> token(1);
	token.memset(0);
	r = ncclAllReduce(token.data(), token.data(), 1, ncclFloat, ncclSum, (ncclComm_t)nccl_comm, m_stream);
	if (r != ncclSuccess) throw std::runtime_error(std::string("ncclAllReduce failed: ") + ncclGetErrorString(r));
	NGP_CUDA_CHECK(cudaStreamSynchronize(m_stream));
}

// in-place all-gather of one state array whose shard [rank * S, (rank + 1) * S) is current on this rank
template <typename T>
static void all_gather_shards(T* base, size_t shard_elems, int rank, ncclDataType_t type, size_t type_scale, void* comm, cudaStream_t stream) {
	const ncclResult_t r = ncclAllGather(base + (size_t)rank * shard_elems, base, shard_elems * type_scale, type, (ncclComm_t)comm, stream);
	if (r != ncclSuccess) throw std::runtime_error(std::string("ncclAllGather failed: ") + ncclGetErrorString(r));
}

void Testbed::gather_optimizer_state() {
	if (!m_peer_enabled) return;
	const size_t S = m_trainer->shard_size();
	all_gather_shards(m_trainer->params_full_precision(), S, m_dp_rank, ncclFloat, 1, m_nccl_comm, m_stream);
	all_gather_shards(m_trainer->m1(), S, m_dp_rank, ncclFloat, 1, m_nccl_comm, m_stream);
	all_gather_shards(m_trainer->m2(), S, m_dp_rank, ncclFloat, 1, m_nccl_comm, m_stream);
	all_gather_shards(m_trainer->param_steps(), S, m_dp_rank, ncclUint32, 1, m_nccl_comm, m_stream);
	NGP_CUDA_CHECK(cudaStreamSynchronize(m_stream));
	m_sharded_state_dirty = false;
}

} // namespace ngp
