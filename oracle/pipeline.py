"""TEST INFRASTRUCTURE ONLY -- the reference's training step (internal/testbed.cu:301-418) executed on the host cores by
chaining the oracle's restatements of its kernels.  Used by bench.py's cpu_baseline / --impl reference legs (timed) and by
tests/ (as the checker of the whole step).  Never imported by the product.  Pinned to the reference's CUDA build through tests/golden/ (see oracle_common.h)."""
import ctypes as C
import time

import numpy as np

from . import (P, u32, f32, u64, i32, lib, Pcg32, trainer_rng, GridLevels, SamplerCfg, hash_encode_fwd, hash_encode_bwd, network_fwd, density_fwd,
               network_fwd_bwd, generate_training_samples, composite_train, rollover, init_params, Optimizer, grid_to_bitfield, grid_mean,
               generate_grid_samples, update_grid, mark_untrained_grid)

CAMERA_ANGLE_X = 0.6911112


def opengl_to_ngp(ogl, scale=0.33, offset=(0.5, 0.5, 0.5)):
    """spec_opengl_to_opencv (internal/utils/common_device.cu:16-30) on [n][3][4] matrices -> [n][12] column-major"""
    m = np.array(ogl, np.float32).copy()
    m[:, :, 1] *= -1
    m[:, :, 2] *= -1
    m[:, :, 3] = m[:, :, 3] * np.float32(scale) + np.array(offset, np.float32)
    m = m[:, [1, 2, 0], :]
    return np.ascontiguousarray(m.transpose(0, 2, 1).reshape(-1, 12))


def sphere_cameras(n, seed=2024, radius=4.0311):
    """OpenGL camera-to-world matrices [n][3][4] on the upper hemisphere (same rig as host/dataset.cpp load_procedural)"""
    rng = Pcg32(seed)
    out = np.zeros((n, 3, 4), np.float32)
    for i in range(n):
        theta = np.float32(2.39996323) * np.float32(i) + np.float32(0.1) * np.float32(rng.next_float())
        phi = np.arcsin(np.float32(0.05) + np.float32(0.9) * (np.float32(i) + np.float32(0.5)) / np.float32(n))
        c = np.array([radius * np.cos(phi) * np.cos(theta), radius * np.cos(phi) * np.sin(theta), radius * np.sin(phi)], np.float64)
        f = -c / np.linalg.norm(c)
        r = np.cross(f, [0, 0, 1.0]); r /= np.linalg.norm(r)
        u = np.cross(r, f)
        out[i, :, 0], out[i, :, 1], out[i, :, 2], out[i, :, 3] = r, u, -f, c
    return out


def render_procedural_views(seed, xforms_ngp, W, H, fx):
    out = np.zeros((xforms_ngp.shape[0], H, W, 4), np.uint8)
    lib().orc_render_procedural_views(u32(seed), u32(xforms_ngp.shape[0]), i32(W), i32(H), f32(fx), f32(fx), P(xforms_ngp), P(out))
    return out


def from_rgba32(rgba8):
    out = np.zeros(rgba8.shape, np.uint16)
    lib().orc_from_rgba32(u64(rgba8.size // 4), P(np.ascontiguousarray(rgba8)), P(out))
    return out


def scene_density_grid(seed, cfg):
    grid = np.zeros(cfg.cells * cfg.n_cascades, np.float32)
    lib().orc_scene_density_grid(u32(seed), u32(cfg.grid_size), u32(cfg.n_cascades), f32(cfg.min_step), P(grid))
    return grid


class CpuTrainer:
    """Testbed::train_nerf_step on the CPU: march -> network -> composite -> rollover -> forward/backward -> Adam/EMA"""

    def __init__(self, n_views=8, res=200, seed=2024, batch=1 << 18, pretrained_occupancy=True):
        self.batch = batch
        self.cfg = SamplerCfg(1)
        self.levels = GridLevels(16, 19, 16)
        self.W = self.H = res
        self.fx = np.float32(0.5 * res / np.tan(0.5 * CAMERA_ANGLE_X))
        self.xforms = opengl_to_ngp(sphere_cameras(n_views, seed))
        self.images = from_rgba32(render_procedural_views(seed, self.xforms, res, res, self.fx))
        rng = trainer_rng(43)
        self.opt = Optimizer(init_params(rng, self.levels.n_params))
        self.rng = Pcg32(43)
        self.sampler_rng = Pcg32(self.rng.next_uint())
        if pretrained_occupancy:
            self.grid = scene_density_grid(seed, self.cfg)
        else:
            self.grid = mark_untrained_grid(self.cfg, self.xforms, res, res, self.fx, self.fx)
        self.mean = grid_mean(self.cfg, self.grid)
        self.bits = grid_to_bitfield(self.cfg, self.grid, self.mean)
        self.n_rays_total = 0
        self.grid_step = 0

    def update_occupancy(self, n_positions):
        """train_prep_nerf (testbed.cu:193-229) on n_positions uniformly chosen cells"""
        pos, idx = generate_grid_samples(self.cfg, n_positions, self.sampler_rng, self.grid_step, self.grid, -0.01)
        self.sampler_rng.advance()
        self.sampler_rng.advance()
        enc = hash_encode_fwd(pos, 3, n_positions, self.opt.w16[10240:].view(np.uint16), self.levels)
        den = density_fwd(self.opt.w16[:10240], enc, n_positions)
        update_grid(self.cfg, idx, den, 16, 0.95, self.grid)
        self.grid_step += 1
        self.mean = grid_mean(self.cfg, self.grid)
        self.bits = grid_to_bitfield(self.cfg, self.grid, self.mean)

    def compute_gradients(self, n_rays, rank=0, world=1):
        """march -> network -> composite -> rollover -> forward/backward of one step; returns (fp16 gradient arena, marched, compacted, loss).
        Data parallel (testbed_train.cpp march_step / learn_step): rank r draws its rays 2^18 * 8 generator draws further along and
        carries 1/world of the loss scale, so the sum of the ranks' gradients is the gradient of the mean loss."""
        w16 = self.opt.w16
        ray_rng = Pcg32(state=self.rng.state, inc=self.rng.inc)
        if rank:
            ray_rng.advance(rank * (1 << 18) * 8)
        s = generate_training_samples(self.cfg, n_rays, self.n_rays_total, self.batch * 16, self.xforms, self.images, self.W, self.H, self.fx, self.fx,
                                      self.bits, ray_rng)
        kept, total = int(s["counters"][0]), int(min(s["counters"][1], self.batch * 16))
        if kept == 0:
            raise RuntimeError("Training generated 0 samples. Aborting training.")
        coords = s["coords"][:total]
        table = w16[10240:].view(np.uint16)
        enc = hash_encode_fwd(coords, 7, total, table, self.levels)
        rgbsigma = network_fwd(w16[:10240], enc, coords, total)
        n_target = min(self.batch, max(total, 1))
        c = composite_train(n_rays, kept, n_target, s["rgba"], s["numsteps"], coords, rgbsigma, self.cfg, self.mean, self.rng, loss_scale=128.0 / world)
        n_comp = min(c["counter"], n_target)
        cc, dl = c["coords_compacted"], c["dloss"]
        if n_comp < n_target:
            dl = rollover(n_target, n_comp, dl, cc)
        enc2 = hash_encode_fwd(cc, 7, n_target, table, self.levels)
        wgrad, denc, _ = network_fwd_bwd(w16[:10240], enc2, cc, dl, n_target)
        ggrid = hash_encode_bwd(cc, 7, n_target, denc, self.levels)
        return np.concatenate([wgrad, ggrid]).astype(np.float16), total, n_comp, float(c["loss"].sum())

    def apply_gradients(self, grads, n_rays):
        self.n_rays_total += n_rays
        self.rng.advance()
        self.opt.step(grads)

    def step(self, n_rays):
        """one training step on n_rays rays; returns (samples marched, samples trained on)"""
        grads, total, n_comp, loss = self.compute_gradients(n_rays)
        self.apply_gradients(grads, n_rays)
        return total, n_comp, loss


def time_cpu_steps(n_steps, n_rays, warmup=1, **kw):
    """samples/s (post-compaction samples per second, the metric of BASELINE.json) of the CPU pipeline"""
    t = CpuTrainer(**kw)
    for _ in range(warmup):
        t.step(n_rays)
    samples, marched, t0 = 0, 0, time.perf_counter()
    for _ in range(n_steps):
        m, n, _ = t.step(n_rays)
        samples += n
        marched += m
    dt = time.perf_counter() - t0
    return dict(samples_per_s=samples / dt, seconds=dt, steps=n_steps, samples_per_step=samples / n_steps, marched_per_step=marched / n_steps, n_rays=n_rays)
