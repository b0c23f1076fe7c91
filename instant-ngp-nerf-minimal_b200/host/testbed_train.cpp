// Testbed::train and its three stages (reference: internal/testbed.cu:168-418).
#include "testbed.h"
#include <nccl.h>
#include <algorithm>
#include <chrono>

namespace ngp {

// testbed.cu:168-191.  The reference brackets each stage with a stream synchronise and a wall clock; here the two stages
// are timed with CUDA events on the stream and the host waits once, at the end, for the counters Python reads.
void Testbed::train(uint32_t n_training_steps, uint32_t target_batch_size) {
	if (!m_trainer || m_training_buffer.dataset.n_images == 0) throw std::runtime_error("train: load_model_config and load_training_data first");
	if (m_dp_stale) throw std::runtime_error("train: load_model_config / load_snapshot rebuilt the model after set_data_parallel; call set_data_parallel again on every rank");
	if (n_training_steps == 0) return;
	m_enqueue_start = std::chrono::steady_clock::now();
	NGP_CUDA_CHECK(cudaEventRecord(m_ev[0], m_stream));
	train_prep_nerf(target_batch_size, n_training_steps, m_stream);
	NGP_CUDA_CHECK(cudaEventRecord(m_ev[1], m_stream));
	train_nerf(target_batch_size, n_training_steps, m_stream);
}

// testbed.cu:193-229.  Data parallel (world > 1): every rank holds the same weights and the same density grid, so the refresh is
// SHARDED instead of replicated -- rank r generates, evaluates and splats only its contiguous 1/world of the sample slots, the
// partial maxima (density_grid_current, non-negative floats) are combined with one ncclAllReduce(max), and the decay + bitfield
// run everywhere on identical input.  max is exact and order-free: the grid is bit-identical to the replicated refresh.
void Testbed::train_prep_nerf(uint32_t, uint32_t, cudaStream_t stream) {
	const uint32_t n_total_elements = m_sampler->n_total_elements;
	const uint32_t n_uniform_positions = m_training_buffer.i_step >= 256 ? n_total_elements / 4 : n_total_elements;
	const uint32_t n_nonuniform_positions = m_training_buffer.i_step >= 256 ? n_total_elements / 4 : 0;
	const uint32_t n_total_positions = n_uniform_positions + n_nonuniform_positions;
	const uint32_t padded_output_width = m_network->padded_density_output_width();
	m_training_buffer.rgbsigma.enlarge((size_t)padded_output_width * n_total_positions);
	if (m_dp_world > 1 && m_shard_grid_refresh) {
		SlotRange ranges[2];
		m_sampler->sample_positions_from_grid(n_uniform_positions, n_nonuniform_positions, m_training_buffer.grid_positions, m_training_buffer.grid_indices, stream,
		                                      (uint32_t)m_dp_rank, (uint32_t)m_dp_world, ranges);
		for (const SlotRange& r : ranges)
			if (r.end > r.begin)
				m_network->density(stream, (const float*)(m_training_buffer.grid_positions.data() + r.begin), sizeof(Position) / sizeof(float), r.end - r.begin,
				                   m_training_buffer.rgbsigma.data() + (size_t)r.begin * padded_output_width, false);
		m_sampler->splat_grid(ranges, 2, padded_output_width, m_training_buffer.rgbsigma, m_training_buffer.grid_indices, m_density_activation, stream);
		const ncclResult_t rc = ncclAllReduce(m_sampler->density_grid_current.data(), m_sampler->density_grid_current.data(), n_total_elements, ncclFloat, ncclMax,
		                                      (ncclComm_t)m_nccl_comm, stream);
		if (rc != ncclSuccess) throw std::runtime_error(std::string("ncclAllReduce(max) of the density grid failed: ") + ncclGetErrorString(rc));
		m_sampler->ema_grid(stream);
	} else {
		m_sampler->sample_positions_from_grid(n_uniform_positions, n_nonuniform_positions, m_training_buffer.grid_positions, m_training_buffer.grid_indices, stream);
		m_network->density(stream, (const float*)m_training_buffer.grid_positions.data(), sizeof(Position) / sizeof(float), n_total_positions,
		                   m_training_buffer.rgbsigma.data(), false);
		m_sampler->update_grid(n_total_positions, padded_output_width, m_training_buffer.rgbsigma, m_training_buffer.grid_indices, m_density_activation, stream);
	}
	m_sampler->update_grid_bitfield(stream);
}

// testbed.cu:231-299
void Testbed::train_nerf(uint32_t target_batch_size, uint32_t n_training_steps, cudaStream_t stream) {
	uint32_t& n_rays = m_training_buffer.n_rays_per_batch;
	TrainingBuffer& tb = m_training_buffer;

	// envmap gradient initialisation (testbed.cu:235-239): the gradients of all steps of this call accumulate, one optimizer step at its end
	if (m_train_envmap) {
		m_envmap.gradients.memset(0, stream);
		m_render->set_envmap(m_envmap.params.data(), m_envmap.gradients.data(), m_envmap.resolution[0], m_envmap.resolution[1], m_envmap.loss_type);
	} else {
		m_render->set_envmap(nullptr, nullptr, 0, 0, ELossType::L2);
	}
	tb.n_coords_counter.enlarge((size_t)n_training_steps * 2);
	NGP_CUDA_CHECK(cudaMemsetAsync(tb.n_coords_counter.data(), 0, sizeof(uint32_t) * 2 * n_training_steps, stream));
	tb.n_compacted_coords_counter.enlarge(n_training_steps);
	NGP_CUDA_CHECK(cudaMemsetAsync(tb.n_compacted_coords_counter.data(), 0, sizeof(uint32_t) * n_training_steps, stream));
	tb.loss.enlarge((size_t)n_rays * n_training_steps);
	NGP_CUDA_CHECK(cudaMemsetAsync(tb.loss.data(), 0, sizeof(float) * n_rays * n_training_steps, stream));

	// Buffers grow before anything is enqueued (a reallocation would synchronise the device in the middle of the pipeline).
	const uint32_t max_samples = target_batch_size * 16;
	for (auto& slot : tb.march) {
		slot.rays.enlarge(n_rays); slot.rays_rgba.enlarge((size_t)n_rays * 4); slot.ray_coord_indices.enlarge((size_t)n_rays * 2); slot.coords.enlarge(max_samples);
	}
	tb.coords_compacted.enlarge(target_batch_size);
	tb.compacted_src_index.enlarge(target_batch_size);
	tb.rgbsigma.enlarge((size_t)max_samples * m_network->padded_output_width());
	tb.dloss_drgbsigma.enlarge((size_t)target_batch_size * m_network->padded_output_width());

	// Software pipeline over the steps of this call.  The march of step i needs only the bitfield (fixed for the whole call) and
	// host-side generator state, so it is issued one step ahead on m_march_stream and hides under step i-1's network and
	// optimizer kernels; the main stream picks its output up through an event.  Results are identical to the serial schedule.
	if (tb.i_step == 0) tb.n_rays_total = 0;
	cudaStream_t march_stream = m_pipeline_march ? m_march_stream : stream;
	NGP_CUDA_CHECK(cudaEventRecord(m_ev_march_go, stream));               // after the refresh and the counter memsets
	NGP_CUDA_CHECK(cudaStreamWaitEvent(march_stream, m_ev_march_go, 0));
	default_rng_t rng_step = m_rng;
	auto issue_march = [&](uint32_t i) {
		const uint32_t slot = i & 1;
		default_rng_t r = m_rng;
		for (uint32_t k = 0; k < i; ++k) r.advance();
		if (i >= 2) NGP_CUDA_CHECK(cudaStreamWaitEvent(march_stream, m_ev_slot_free[slot], 0));   // step i-2 has consumed this slot
		march_step(slot, target_batch_size, n_rays, tb.n_rays_total + n_rays * i, r, tb.n_coords_counter.data() + 2 * i, march_stream);
		NGP_CUDA_CHECK(cudaEventRecord(m_ev_march_ready[slot], march_stream));
	};
	issue_march(0);
	for (uint32_t i = 0; i < n_training_steps; ++i) {
		const uint32_t slot = i & 1;
		if (i + 1 < n_training_steps) issue_march(i + 1);
		NGP_CUDA_CHECK(cudaStreamWaitEvent(stream, m_ev_march_ready[slot], 0));
		learn_step(slot, target_batch_size, n_rays, rng_step, tb.n_coords_counter.data() + 2 * i, tb.n_compacted_coords_counter.data() + i,
		           tb.loss.data() + (size_t)n_rays * i, stream);
		NGP_CUDA_CHECK(cudaEventRecord(m_ev_slot_free[slot], stream));
		rng_step.advance();
		++tb.i_step;
	}
	m_rng = rng_step;
	tb.n_rays_total += n_rays * n_training_steps;

	if (m_peer_enabled && m_trainer->ema() != m_trainer->params()) {
		// (measured at 4 GPUs: skipping this all-gather shortens the call boundary by 90 us = 6 us per step)
		// the EMA (inference) weights are sharded like the rest of the optimizer state: collect them once per call, so render() and
		// plain snapshots stay local operations
		const size_t S = m_trainer->shard_size();
		const ncclResult_t r = ncclAllGather(m_trainer->ema() + (size_t)m_dp_rank * S, m_trainer->ema(), S, ncclHalf, (ncclComm_t)m_nccl_comm, stream);
		if (r != ncclSuccess) throw std::runtime_error(std::string("ncclAllGather failed: ") + ncclGetErrorString(r));
	}
	if (m_train_envmap) envmap_optimizer_step(stream, LOSS_SCALE * (float)n_training_steps);      // testbed.cu:295-298
	api_check(ngp_reduce_sum(stream, tb.loss.data(), (uint64_t)n_rays * n_training_steps, tb.loss_sum.data()));
	NGP_CUDA_CHECK(cudaEventRecord(m_ev[2], stream));

	tb.enqueue_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - m_enqueue_start).count();
	// GPU -> CPU: the one synchronisation of a train() call
	std::vector<uint32_t> pairs((size_t)n_training_steps * 2);
	tb.n_compacted_coords_counter_cpu.resize(n_training_steps);
	NGP_CUDA_CHECK(cudaMemcpyAsync(pairs.data(), tb.n_coords_counter.data(), pairs.size() * 4, cudaMemcpyDeviceToHost, stream));
	NGP_CUDA_CHECK(cudaMemcpyAsync(tb.n_compacted_coords_counter_cpu.data(), tb.n_compacted_coords_counter.data(), n_training_steps * 4, cudaMemcpyDeviceToHost, stream));
	float loss_sum = 0.f;
	NGP_CUDA_CHECK(cudaMemcpyAsync(&loss_sum, tb.loss_sum.data(), 4, cudaMemcpyDeviceToHost, stream));
	NGP_CUDA_CHECK(cudaStreamSynchronize(stream));
	NGP_CUDA_CHECK(cudaEventElapsedTime(&tb.training_prep_ms, m_ev[0], m_ev[1]));
	NGP_CUDA_CHECK(cudaEventElapsedTime(&tb.training_ms, m_ev[1], m_ev[2]));

	tb.n_coords_counter_cpu.resize(n_training_steps);
	tb.measured_batch_size_before_compaction = 0;
	tb.measured_batch_size = 0;
	uint64_t before = 0, after = 0;
	for (uint32_t i = 0; i < n_training_steps; ++i) {
		tb.n_coords_counter_cpu[i] = pairs[2 * i + 1];
		if (tb.n_coords_counter_cpu[i] == 0 || tb.n_compacted_coords_counter_cpu[i] == 0) throw std::runtime_error("Training generated 0 samples. Aborting training.");
		before += tb.n_coords_counter_cpu[i];
		after += tb.n_compacted_coords_counter_cpu[i];
	}
	tb.measured_batch_size_before_compaction = (uint32_t)(before / n_training_steps);
	tb.measured_batch_size = (uint32_t)(after / n_training_steps);

	tb.loss_scalar = loss_sum / (float)n_training_steps;
	tb.loss_scalar *= (float)tb.measured_batch_size / (float)target_batch_size;

	// dynamically determine the number of rays for each step (testbed.cu:291-293)
	n_rays = (uint32_t)((float)n_rays * (float)target_batch_size / (float)tb.measured_batch_size);
	n_rays = std::min(next_multiple(n_rays, BATCH_SIZE_MULTIPLE), 1u << 18);
}

// testbed.cu:301-344: rays + occupancy march + ordered compaction of one step, into march slot `slot`
void Testbed::march_step(uint32_t slot, uint32_t target_batch_size, uint32_t n_rays_per_batch, uint32_t n_rays_total, default_rng_t rng, uint32_t* counters,
                         cudaStream_t stream) {
	TrainingBuffer& tb = m_training_buffer;
	TrainingBuffer::MarchSlot& ms = tb.march[slot];
	const uint32_t max_samples = target_batch_size * 16;
	const uint32_t max_inference = next_multiple(std::min(tb.measured_batch_size_before_compaction, max_samples), BATCH_SIZE_MULTIPLE);
	// each data-parallel rank draws its own pixels: the per-ray generator is advanced by i*8 draws inside the kernel, so rank r
	// starts 2^18 * 8 draws further along (2^18 = the cap on rays per batch)
	default_rng_t ray_rng = rng;
	if (m_dp_rank) ray_rng.advance((int64_t)m_dp_rank * (1ll << 18) * 8);
	m_sampler->sample_coords_from_dataset(n_rays_per_batch, n_rays_total, max_inference, tb.dataset, m_train_envmap, ms.rays, ms.rays_rgba, counters,
	                                      ms.ray_coord_indices, ms.coords, counters + 1, ray_rng, stream);
}

// testbed.cu:346-418: everything of one step that depends on the current weights
void Testbed::learn_step(uint32_t slot, uint32_t target_batch_size, uint32_t n_rays_per_batch, default_rng_t rng, uint32_t* counters, uint32_t* compacted_counter,
                         float* loss, cudaStream_t stream) {
	TrainingBuffer& tb = m_training_buffer;
	TrainingBuffer::MarchSlot& ms = tb.march[slot];
	const uint32_t max_samples = target_batch_size * 16;
	const uint32_t padded_output_width = m_network->padded_output_width();
	const uint32_t max_inference = next_multiple(std::min(tb.measured_batch_size_before_compaction, max_samples), BATCH_SIZE_MULTIPLE);

	// 2. network on every marched sample (training weights); the device-side count bounds the work
	m_network->inference_mixed_precision(stream, ms.coords.data(), max_inference, counters + 1, tb.rgbsigma.data(), false);
	// 3. compositing, loss, dL/d(rgb sigma), compaction.  Data parallel: every rank's gradient carries 1/world of the loss scale,
	//    so the NCCL sum below is the mean over ranks and stays inside fp16 range.
	m_render->volume_rendering_with_loss_and_gradient(
		n_rays_per_batch, target_batch_size, ms.rays, ms.rays_rgba, counters, ms.ray_coord_indices, ms.coords, tb.coords_compacted, compacted_counter,
		tb.rgbsigma, tb.dloss_drgbsigma, (int)padded_output_width, m_rgb_activation, m_density_activation, loss, m_loss_type, LOSS_SCALE / (float)m_dp_world,
		m_sampler->aabb, m_sampler->n_cascades, m_sampler->min_cone_stepsize, m_sampler->density_grid_mean, m_sampler->min_optical_thickness,
		m_background_color, m_color_space, rng, stream, tb.compacted_src_index.data());
	// 4. rollover
	api_check(ngp_rollover(stream, target_batch_size, compacted_counter, tb.dloss_drgbsigma.data(), padded_output_width, (float*)tb.coords_compacted.data(),
	                       tb.compacted_src_index.data()));
	// 5./6. forward + backward on the compacted batch (one fused launch chain).  The compacted samples are a subset of the ones
	//       step 2 encoded with the same table, so their encodings are gathered from there instead of being computed again.
	m_network->forward(stream, tb.coords_compacted.data(), target_batch_size, tb.rgbsigma.data());
	m_network->backward(stream, tb.coords_compacted.data(), target_batch_size, tb.rgbsigma.data(), tb.dloss_drgbsigma.data(),
	                    tb.compacted_src_index.data(), max_inference);
	if (m_peer_enabled) {
		// data parallel over peer memory: publish "my gradients are complete", then ONE kernel reduces this rank's shard out of all
		// ranks' arenas, updates it and stores the new weights into all arenas; the next reader of the weights waits for every shard
		++m_peer_epoch;
		m_sharded_state_dirty = true;
		api_check(ngp_peer_publish_gradients(stream, &m_peers, m_peer_epoch));
		m_trainer->optimizer_step_sharded(stream, LOSS_SCALE, m_peers, m_peer_epoch);
		api_check(ngp_peer_wait_weights(stream, &m_peers, m_peer_epoch));
		return;
	}
	// data parallel through NCCL: sum the fp16 gradient arena over ranks (NVLink / NVSwitch), every rank runs the whole optimizer
	if (m_dp_world > 1) {
		const ncclResult_t r = ncclAllReduce(m_trainer->gradients(), m_trainer->gradients(), m_trainer->n_params(), ncclHalf, ncclSum, (ncclComm_t)m_nccl_comm, stream);
		if (r != ncclSuccess) throw std::runtime_error(std::string("ncclAllReduce failed: ") + ncclGetErrorString(r));
	}
	// 7. optimizer
	m_trainer->optimizer_step(stream, LOSS_SCALE);
}

// the reference's single-step entry point (testbed.cu:301-418), kept for callers that drive steps themselves: serial schedule
void Testbed::train_nerf_step(uint32_t target_batch_size, uint32_t n_rays_per_batch, uint32_t* counters, uint32_t* compacted_counter, float* loss, cudaStream_t stream) {
	TrainingBuffer& tb = m_training_buffer;
	const uint32_t max_samples = target_batch_size * 16;
	TrainingBuffer::MarchSlot& ms = tb.march[0];
	ms.rays.enlarge(n_rays_per_batch); ms.rays_rgba.enlarge((size_t)n_rays_per_batch * 4); ms.ray_coord_indices.enlarge((size_t)n_rays_per_batch * 2);
	ms.coords.enlarge(max_samples);
	tb.coords_compacted.enlarge(target_batch_size);
	tb.compacted_src_index.enlarge(target_batch_size);
	tb.rgbsigma.enlarge((size_t)max_samples * m_network->padded_output_width());
	tb.dloss_drgbsigma.enlarge((size_t)target_batch_size * m_network->padded_output_width());
	if (tb.i_step == 0) tb.n_rays_total = 0;
	march_step(0, target_batch_size, n_rays_per_batch, tb.n_rays_total, m_rng, counters, stream);
	tb.n_rays_total += n_rays_per_batch;
	learn_step(0, target_batch_size, n_rays_per_batch, m_rng, counters, compacted_counter, loss, stream);
	m_rng.advance();
	++tb.i_step;
}

} // namespace ngp
