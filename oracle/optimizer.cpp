// TEST INFRASTRUCTURE ONLY (see oracle_common.h).  Pinned to golden vectors from the reference's CUDA build (tests/golden/).
// CPU restatement of parameter initialisation and the optimizer chain:
//   Trainer ctor / initialize_params : tiny-cuda-nn/trainer.h:54-112
//   xavier uniform                   : tiny-cuda-nn/gpu_matrix.h:147-164
//   grid init (generate_random_uniform, 4 values / thread, strided) : random.h:69-94, encodings/grid.h:720-733
//   Adam                             : optimizers/adam.h:50-113
//   EMA (half precision variant)     : optimizers/ema.h:65-79, 104-136
//   exponential decay                : optimizers/exponential_decay.h:62-73
#include "oracle_common.h"
#include <random>
#include <omp.h>

using namespace orc;

extern "C" {

// seed as used by Trainer: std::seed_seq{seed}.generate(2 values), pcg32{first}
void orc_trainer_rng(uint32_t seed, uint64_t* state, uint64_t* inc) {
	std::seed_seq seq{seed};
	std::vector<uint32_t> seeds(2);
	seq.generate(seeds.begin(), seeds.end());
	Pcg32 r{(uint64_t)seeds.front()};
	*state = r.state; *inc = r.inc;
}

// generic helpers so tests can drive a pcg32 stream
void orc_pcg32_seed(uint64_t initstate, uint64_t initseq, uint64_t* state, uint64_t* inc) { Pcg32 r{initstate, initseq}; *state = r.state; *inc = r.inc; }
void orc_pcg32_advance(uint64_t* state, uint64_t inc, int64_t delta) { Pcg32 r; r.state = *state; r.inc = inc; r.advance(delta); *state = r.state; }
uint32_t orc_pcg32_next_uint(uint64_t* state, uint64_t inc) { Pcg32 r; r.state = *state; r.inc = inc; uint32_t v = r.next_uint(); *state = r.state; return v; }
float orc_pcg32_next_float(uint64_t* state, uint64_t inc) { Pcg32 r; r.state = *state; r.inc = inc; float v = r.next_float(); *state = r.state; return v; }

// params_fp32 [n_mlp + n_grid]: MLP matrices xavier-uniform drawn one float at a time on the host, then the grid
// U(-1e-4,1e-4) with the strided 4-per-thread pattern of generate_random_kernel (threads rounded up to 128).
// layer_shapes: n_layers x (rows, cols).  The generator state is advanced exactly as the reference does.
void orc_init_params(uint64_t* state, uint64_t inc, const uint32_t* layer_shapes, uint32_t n_layers, uint64_t n_grid, float* params) {
	Pcg32 rnd; rnd.state = *state; rnd.inc = inc;
	size_t pos = 0;
	for (uint32_t l = 0; l < n_layers; ++l) {
		uint32_t rows = layer_shapes[l * 2], cols = layer_shapes[l * 2 + 1];
		float scale = std::sqrt(6.0f / (float)(rows + cols));
		for (size_t i = 0; i < (size_t)rows * cols; ++i) params[pos++] = rnd.next_float() * 2.0f * scale - scale;
	}
	const uint64_t n_threads = (n_grid + 3) / 4;
	const uint64_t total_threads = ((n_threads + 127) / 128) * 128;
	float* grid = params + pos;
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)total_threads; ++i) {      // every launched thread writes, also those past div_round_up(n, 4) (random.h:69-84)
		Pcg32 r = rnd;
		r.advance((int64_t)i * 4);
		for (uint64_t j = 0; j < 4; ++j) {
			uint64_t idx = (uint64_t)i + total_threads * j;
			if (idx >= n_grid) break;
			grid[idx] = fmaf(r.next_float(), 1e-4f - (-1e-4f), -1e-4f);   // val * (upper - lower) + lower, contracted by nvcc (random.h:99)
		}
	}
	rnd.advance((int64_t)n_grid);
	*state = rnd.state;
}

// One optimizer step over the arena.  hyper: lr_base, beta1, beta2, eps, l2_reg, loss_scale, ema_decay,
// decay_base; sched: decay_start, decay_interval.  state_io: [0]=adam step count (uint), [1]=lr factor bits.
void orc_optimizer_step(uint64_t n, uint64_t n_matrix, const float* hyper, const uint32_t* sched, uint32_t* step_io, float* lr_factor_io,
                        float* w32, uint16_t* w16, const uint16_t* g16, float* m1, float* m2, uint32_t* psteps, uint16_t* ema16) {
	const float lr_base = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], l2 = hyper[4], loss_scale = hyper[5], ema_decay = hyper[6], decay_base = hyper[7];
	uint32_t step = *step_io;
	if (step == 0) *lr_factor_io = 1.0f;
	if (step >= sched[0] && (step - sched[0]) % sched[1] == 0) *lr_factor_io *= decay_base;
	const float lr0 = lr_base * *lr_factor_io;
	++step;
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)n; ++i) {
		float g = h16from(g16[i]) / loss_scale;
		const bool matrix = (uint64_t)i < n_matrix;
		if (!(!matrix && g == 0)) {
			const float w = w32[i];
			if (matrix) g += l2 * w;
			const float gsq = g * g;
			const float fm = m1[i] = b1 * m1[i] + (1 - b1) * g;
			const float sm = m2[i] = b2 * m2[i] + (1 - b2) * gsq;
			const uint32_t cs = ++psteps[i];
			float lr = lr0 * (sqrtf(1 - powf(b2, (float)cs)) / (1 - powf(b1, (float)cs)));
			const float eff = fminf(fmaxf(lr / (sqrtf(sm) + eps), 0.f), 3.402823466e38f);
			const float nw = w - eff * fm;
			w32[i] = nw;
			w16[i] = h16bits(nw);
		}
	}
	const float deb_old = 1 - (float)std::pow(ema_decay, step - 1);
	const float deb_new = 1.0f / (1 - (float)std::pow(ema_decay, step));
	#pragma omp parallel for schedule(static)
	for (int64_t i = 0; i < (int64_t)n; ++i) {
		float f = (h16from(ema16[i]) * ema_decay * deb_old + h16from(w16[i]) * (1 - ema_decay)) * deb_new;
		ema16[i] = h16bits(f);
	}
	*step_io = step;
}

} // extern "C"
