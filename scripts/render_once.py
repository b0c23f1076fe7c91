#!/usr/bin/env python
"""Two-stage helper for profiling the render path alone.
  python scripts/render_once.py train <snapshot> [calls=40]     train, save a snapshot
  python scripts/render_once.py render <snapshot> [W H frames]  load it, render `frames` frames (run THIS stage under ncu)"""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pkg = importlib.import_module("instant-ngp-nerf-minimal_b200")
pkg.capi.load()
pyngp = pkg.load_pyngp()


def main():
    stage, snap = sys.argv[1], sys.argv[2]
    tb = pyngp.Testbed()
    if stage == "train":
        calls = int(sys.argv[3]) if len(sys.argv) > 3 else 40
        tb.load_model_config(os.path.join(ROOT, "instant-ngp-nerf-minimal_b200", "configs", "base.json"))
        tb.load_procedural_training_data(100, 800, 800, 2024, 0.33, [0.5, 0.5, 0.5])
        for _ in range(calls):
            tb.train(16, 1 << 18)
        tb.save_snapshot(snap, False)
        return
    w, h, frames = (int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])) if len(sys.argv) > 5 else (1920, 1080, 1)
    tb.load_snapshot(snap)
    tb.load_procedural_training_data(100, 800, 800, 2024, 0.33, [0.5, 0.5, 0.5])      # cameras only (view 3 below)
    tb.load_snapshot(snap)
    tb.set_camera_to_training_view(3)
    tb.background_color = [0, 0, 0, 1]
    if os.environ.get("NGP_BANDS"):
        tb.rendering_buffer.CONCURRENT_BANDS = int(os.environ["NGP_BANDS"])
    if os.environ.get("NGP_ROWS"):
        r0, r1 = (int(x) for x in os.environ["NGP_ROWS"].split(":"))
        tb.set_render_rows(r0, r1)
    for _ in range(frames):
        t0 = time.perf_counter()
        ms = tb.render_frames_ms(w, h, 1, True)
        rb = tb.rendering_buffer
        print(f"{w}x{h}: {ms:.2f} ms ({time.perf_counter() - t0:.4f} s wall), {rb.rounds_last_frame} rounds, {rb.readbacks_last_frame} read-backs")


if __name__ == "__main__":
    main()
