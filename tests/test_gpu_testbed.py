"""GPU tests through the reference's own surface: pyngp.Testbed (internal/python_api.cu:48-139) -- load_model_config,
load data, train(n_steps, batch), render(w, h, spp, to_srgb), save / load_snapshot -- on a small procedural scene.

What is checked here is what only the assembled path can show: the loss goes down, the counters Python reads are filled, a
snapshot reproduces the rendered image bit for bit, a tile-sharded render (BASELINE configs[3]: row bands, one per GPU, no
collective) is bit-identical to the full frame, and the large-table cascaded configuration (configs[4]: T = 2^22, aabb_scale
16 -> 5 cascades) trains.  Kernel-level parity against the oracle / the reference's CUDA build lives in the other test files.
"""
import importlib
import json
import os
import re

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "instant-ngp-nerf-minimal_b200"
BASE = os.path.join(ROOT, PKG, "configs", "base.json")


@pytest.fixture(scope="module")
def pyngp():
    pkg = importlib.import_module(PKG)
    pkg.capi.load()                      # raises when the CUDA library is missing: no fallback
    return pkg.load_pyngp()


def _config(tmp_path, **edits):
    """base.json with some values replaced (the file carries // comments, like the reference's json reader accepts)"""
    text = re.sub(r"^\s*//.*$", "", open(BASE).read(), flags=re.M)
    cfg = json.loads(text)
    for path, value in edits.items():
        node = cfg
        keys = path.split(".")
        for k in keys[:-1]:
            node = node[k]
        node[keys[-1]] = value
    out = os.path.join(tmp_path, "config.json")
    json.dump(cfg, open(out, "w"))
    return out


def _trained(pyngp, calls=6, batch=1 << 16, config=BASE, views=12, res=160):
    tb = pyngp.Testbed()
    tb.load_model_config(config)
    tb.load_procedural_training_data(views, res, res, 2024, 0.33, [0.5, 0.5, 0.5])
    losses = []
    for _ in range(calls):
        tb.train(16, batch)
        losses.append(tb.training_buffer.loss)
    return tb, losses


def test_train_fills_the_counters_and_the_loss_goes_down(pyngp):
    tb, losses = _trained(pyngp, calls=10)
    b = tb.training_buffer
    assert b.i_step == 160
    assert len(b.n_coords_counter_cpu) == 16 and len(b.n_compacted_coords_counter_cpu) == 16
    assert all(c > 0 for c in b.n_coords_counter_cpu) and all(c > 0 for c in b.n_compacted_coords_counter_cpu)
    assert 0 < b.measured_batch_size <= b.measured_batch_size_before_compaction
    assert b.n_rays_per_batch % 256 == 0 and b.n_rays_per_batch > 0
    assert b.training_ms > 0 and b.training_prep_ms > 0
    assert np.all(np.isfinite(losses))
    assert np.mean(losses[-3:]) < 0.5 * losses[0], losses


def test_train_requires_config_and_data(pyngp):
    tb = pyngp.Testbed()
    with pytest.raises(RuntimeError):
        tb.train(1, 1 << 14)
    with pytest.raises(RuntimeError):
        tb.load_model_config("/nonexistent/config.json")


def test_pipelined_and_serial_march_give_the_same_training(pyngp):
    """the look-ahead march is a schedule, not a different computation: sample counts per step are identical"""
    out = []
    for pipelined in (True, False):
        tb = pyngp.Testbed()
        tb.load_model_config(BASE)
        tb.load_procedural_training_data(12, 160, 160, 2024, 0.33, [0.5, 0.5, 0.5])
        tb.pipeline_march = pipelined
        tb.train(16, 1 << 16)
        out.append((list(tb.training_buffer.n_coords_counter_cpu), tb.training_buffer.n_rays_per_batch))
    assert out[0][0][0] == out[1][0][0]                 # first step: same weights, same rays -> same march, bit for bit
    assert out[0][1] == out[1][1] or abs(out[0][1] - out[1][1]) <= 256


def test_render_row_bands_are_bit_identical_to_the_full_frame(pyngp):
    tb, _ = _trained(pyngp)
    tb.set_camera_to_training_view(3)
    tb.background_color = [0, 0, 0, 1]
    w, h = 320, 184
    full = np.array(tb.render(w, h, 1, True))
    assert full.shape == (h, w, 4) and np.isfinite(full).all()
    assert full[..., 3].max() > 0.5, "the trained scene must be visible"
    for n_bands in (8, 3):
        edges = [round(i * h / n_bands) for i in range(n_bands + 1)]
        stitched = np.zeros_like(full)
        for r0, r1 in zip(edges[:-1], edges[1:]):
            tb.set_render_rows(r0, r1)
            band = np.array(tb.render(w, h, 1, True))
            stitched[r0:r1] = band[r0:r1]
            assert not band[:r0].any() and not band[r1:].any(), "rows outside the band read zero"
            alone = np.array(tb.render_band(w, h, 1, True))
            assert alone.shape == (r1 - r0, w, 4) and np.array_equal(alone.view(np.uint32), band[r0:r1].view(np.uint32))
        tb.set_render_rows(0, -1)
        assert np.array_equal(stitched.view(np.uint32), full.view(np.uint32)), f"{n_bands} bands differ from the full frame"
    tb.set_render_rows(7, 7)                             # empty band: nothing marched, nothing crashes
    empty = np.array(tb.render(w, h, 1, True))
    tb.set_render_rows(0, -1)
    assert empty.shape == full.shape
    again = np.array(tb.render(w, h, 1, True))
    assert np.array_equal(again.view(np.uint32), full.view(np.uint32)), "rendering is deterministic"


def test_block_skipping_march_renders_the_identical_image(pyngp):
    """Testbed.coarse_skip: the inference march crosses empty 8x8x8 blocks of the occupancy grid in one hop (csrc/march_infer.cu
    coarse_hop); with it and without it every pixel of every render mode is the same bit pattern"""
    tb, _ = _trained(pyngp)
    tb.background_color = [0, 0, 0, 1]
    assert tb.coarse_skip
    for view, (w, h) in [(3, (400, 240)), (11, (256, 256))]:
        tb.set_camera_to_training_view(view)
        for mode in (pyngp.RenderMode.Shade, pyngp.RenderMode.Depth, pyngp.RenderMode.Cost):
            tb.rendering_buffer.render_mode = mode
            tb.coarse_skip = True
            fast = np.array(tb.render(w, h, 1, False))
            tb.coarse_skip = False
            plain = np.array(tb.render(w, h, 1, False))
            tb.coarse_skip = True
            assert np.array_equal(fast.view(np.uint32), plain.view(np.uint32)), (view, mode)
    tb.rendering_buffer.render_mode = pyngp.RenderMode.Shade


def test_device_driven_render_loop_gives_the_identical_image(pyngp):
    """rendering_buffer.device_loop: the rounds of Testbed::render_frame (testbed.cu:521-575) decided by a kernel instead of the host,
    with late rounds taking more steps (TAIL_STEPS).  Against the host loop with the reference's schedule: same bits in every
    pixel, all render modes (Cost shows the per-ray step counter the rounds hand on), far fewer host synchronisations."""
    tb, _ = _trained(pyngp)
    tb.background_color = [0, 0, 0, 1]
    rb = tb.rendering_buffer
    assert rb.device_loop
    rb.CONCURRENT_BANDS_MIN_PIXELS = 0                    # the test frames are small: let the two-stream path run on them
    tb.set_camera_to_training_view(5)
    w, h = 384, 216
    for mode in (pyngp.RenderMode.Shade, pyngp.RenderMode.Normals, pyngp.RenderMode.Depth, pyngp.RenderMode.Cost):
        rb.render_mode = mode
        rb.device_loop = False
        want = np.array(tb.render(w, h, 1, False))
        host_rounds, host_syncs = rb.rounds_last_frame, rb.readbacks_last_frame
        for tail, per_batch, bands, min_batch, fused in ((32, 8, 2, 1 << 21, True), (32, 8, 1, 0, False), (8, 3, 1, 0, True), (64, 1, 2, 1 << 18, False)):
            rb.device_loop, rb.TAIL_STEPS, rb.ROUNDS_PER_READBACK, rb.CONCURRENT_BANDS, rb.MIN_BATCH_SAMPLES = True, tail, per_batch, bands, min_batch
            rb.fused_compaction = fused
            got = np.array(tb.render(w, h, 1, False))
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (mode, tail, per_batch, bands, min_batch)
            if (tail, per_batch) == (32, 8):
                assert rb.readbacks_last_frame * 3 < host_syncs, (rb.readbacks_last_frame, host_syncs)
            if tail == 8:
                assert rb.rounds_last_frame >= host_rounds          # the reference's schedule, plus the empty rounds of the last batch
    rb.render_mode = pyngp.RenderMode.Shade
    rb.device_loop, rb.TAIL_STEPS, rb.ROUNDS_PER_READBACK, rb.CONCURRENT_BANDS, rb.MIN_BATCH_SAMPLES = True, 32, 8, 2, 1 << 21
    rb.CONCURRENT_BANDS_MIN_PIXELS = 1 << 20
    rb.fused_compaction = True
    assert want[..., 3].max() > 0.5


def test_render_modes_and_tonemap_curves(pyngp):
    tb, _ = _trained(pyngp)
    tb.set_camera_to_training_view(0)
    tb.background_color = [0, 0, 0, 1]
    w, h = 160, 160
    shade = np.array(tb.render(w, h, 1, False))
    rb = tb.rendering_buffer
    for curve in (pyngp.TonemapCurve.ACES, pyngp.TonemapCurve.Hable, pyngp.TonemapCurve.Reinhard):
        rb.tonemap_curve = curve
        img = np.array(tb.render(w, h, 1, False))
        assert np.isfinite(img).all() and not np.array_equal(img[..., :3], shade[..., :3])
        assert np.array_equal(img[..., 3], shade[..., 3]), "the curve acts on colour only"
    rb.tonemap_curve = pyngp.TonemapCurve.Reinhard
    img = np.array(tb.render(w, h, 1, False))
    lum = 0.2126 * shade[..., 0] + 0.7152 * shade[..., 1] + 0.0722 * shade[..., 2]
    ref = shade[..., :3] / (lum[..., None] + 1.0)         # utils/render_buffer.cu:131-135 on the linear image
    assert np.allclose(img[..., :3], np.maximum(ref, 0), atol=2e-6)
    rb.tonemap_curve = pyngp.TonemapCurve.Identity
    rb.render_mode = pyngp.RenderMode.Depth
    depth = np.array(tb.render(w, h, 1, False))
    assert np.isfinite(depth).all() and depth[..., 0].max() > 0
    rb.render_mode = pyngp.RenderMode.Normals              # network input gradients (testbed.cu:558-560, ray_marcher.cu:260-266)
    tb.background_color = [0, 0, 0, 0]                     # transparent: alpha is the object's own coverage
    nrm = np.array(tb.render(w, h, 1, False))
    tb.background_color = [0, 0, 0, 1]
    rb.render_mode = pyngp.RenderMode.Shade
    assert np.isfinite(nrm).all()
    hit = nrm[..., 3] > 0.9
    assert hit.sum() > 200, "the object must be visible"
    n = nrm[hit][:, :3] / nrm[hit][:, 3:4] * 2 - 1           # shade: (0.5 n + 0.5) * alpha
    assert np.all(np.abs(np.linalg.norm(n, axis=1) - 1) < 2e-3), "shaded normals are unit vectors"
    assert not np.array_equal(nrm[..., :3], shade[..., :3])


def _load_and_render(pyngp, path):
    tb = pyngp.Testbed()
    tb.load_model_config(BASE)
    tb.load_procedural_training_data(12, 160, 160, 2024, 0.33, [0.5, 0.5, 0.5])
    tb.load_snapshot(path)                                # network, weights and density grid come from the file
    tb.set_camera_to_training_view(2)
    tb.background_color = [0, 0, 0, 1]
    return tb, np.array(tb.render(200, 120, 1, True))


def test_snapshot_round_trip_renders_identically(pyngp, tmp_path):
    tb, _ = _trained(pyngp)
    tb.set_camera_to_training_view(2)
    tb.background_color = [0, 0, 0, 1]
    before = np.array(tb.render(200, 120, 1, True))
    path = os.path.join(tmp_path, "snap.msgpack")
    tb.save_snapshot(path, False)
    assert os.path.getsize(path) > 20 << 20              # 12.2 M fp16 parameters + the density grid
    tb2, after = _load_and_render(pyngp, path)
    # the file carries the weights render() uses or the training weights (testbed.cu:127-141 / trainer.h:294-317): either way the
    # picture is the same scene, and a second trip through a file is exact
    mse = float(np.mean((before[..., :3] - after[..., :3]) ** 2))
    assert mse < 1e-3, mse
    path2 = os.path.join(tmp_path, "snap2.msgpack")
    tb2.save_snapshot(path2, False)
    _, again = _load_and_render(pyngp, path2)
    assert np.array_equal(after.view(np.uint32), again.view(np.uint32))
    with pytest.raises(RuntimeError):
        pyngp.Testbed().load_snapshot(os.path.join(tmp_path, "missing.msgpack"))


def test_trainable_environment_map(pyngp, tmp_path):
    """envmap.train (testbed.cu:82-90, 235-239, 295-298): the map starts at zero, collects gradients over the steps of a train()
    call, steps its own optimizer once per call, and render() starts every pixel from it"""
    cfg = _config(str(tmp_path), **{"envmap.train": True, "envmap.height": 64, "envmap.width": 32, "render.train_with_random_bg_color": False})
    tb = pyngp.Testbed()
    tb.load_model_config(cfg)
    assert not np.array(tb.envmap_image()).any()
    tb.background_color = [0.3, 0.6, 0.2, 1.0]
    tb.load_procedural_training_data(12, 160, 160, 2024, 0.33, [0.5, 0.5, 0.5])
    losses = []
    for _ in range(8):
        tb.train(16, 1 << 16)
        losses.append(tb.training_buffer.loss)
    assert np.all(np.isfinite(losses)) and np.mean(losses[-3:]) < 0.5 * losses[0], losses
    env = np.array(tb.envmap_image())
    assert env.shape == (32, 64, 4) and np.isfinite(env).all()
    assert np.count_nonzero(env[..., :3]) > 100, "rays that reach the background train the map"
    assert not env[..., 3].any(), "alpha receives no gradient (ray_marcher.cu:208-209)"
    tb.set_camera_to_training_view(1)
    img = np.array(tb.render(96, 96, 1, True))
    assert np.isfinite(img).all()
    # the plain configuration is untouched by the feature: its map reads zero and its training is the one of the other tests
    tb0, l0 = _trained(pyngp, calls=2)
    assert not np.array(tb0.envmap_image()).any() and np.all(np.isfinite(l0))


def test_large_table_cascaded_configuration_trains(pyngp, tmp_path):
    """BASELINE configs[4]: log2_hashmap_size 22 (150 MB fp16 table, spills L2), aabb_scale 16 (5 cascades), 2^20-sample batches"""
    cfg = _config(str(tmp_path), **{"samlper.aabb_scale": 16, "network.encoding.log2_hashmap_size": 22})
    tb = pyngp.Testbed()
    tb.load_model_config(cfg)
    assert tb.n_params == 10240 + 78948048               # 39 474 024 entries x 2
    tb.load_procedural_training_data(12, 160, 160, 2024, 0.33, [0.5, 0.5, 0.5])
    losses = []
    for _ in range(5):
        tb.train(16, 1 << 20)
        losses.append(tb.training_buffer.loss)
    b = tb.training_buffer
    assert all(c > 0 for c in b.n_compacted_coords_counter_cpu)
    assert np.all(np.isfinite(losses)) and losses[-1] < 0.7 * losses[0], losses
    tb.set_camera_to_training_view(1)
    img = np.array(tb.render(160, 160, 1, True))
    assert np.isfinite(img).all()
