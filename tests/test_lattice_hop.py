"""CPU: the closed-form empty-voxel hop (csrc/lattice_hop.h, used by the inference march kernels) against the loop it replaces --
the tail of advance_to_next_voxel, internal/sampler/occupancy_sampler.cu:284-294:  do { t += dt; } while (t < t_target);
Bit-exact on every input where the closed form claims to apply (it may decline: binade crossings, ties), on millions of random
(t, target) pairs including targets placed exactly on, one ulp below and one ulp above lattice points."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def shim(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("lattice") / "shim.so")
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-I", os.path.join(ROOT, "instant-ngp-nerf-minimal_b200", "csrc"),
                           os.path.join(ROOT, "tests", "lattice_hop_shim.c"), "-o", so, "-lm"])
    lib = C.CDLL(so)
    lib.run.restype = C.c_long
    lib.run_cached.restype = C.c_long
    lib.loop.restype = C.c_float
    lib.loop.argtypes = [C.c_float] * 3
    lib.hop.argtypes = [C.c_float] * 3 + [C.POINTER(C.c_float)]
    return lib


@pytest.mark.parametrize("dt", [1.73205080757 / 1024, 1.73205080757 / 4096, 2.0 ** -9, 1.5 * 2.0 ** -9, 1.73205080757 / 256])
@pytest.mark.parametrize("tmax", [0.9, 3.0, 30.0])
def test_closed_form_hop_equals_the_loop(shim, dt, tmax):
    used = C.c_long()
    n = 1500000
    bad = shim.run(C.c_long(n), C.c_ulonglong(int(dt * 1e9) + int(tmax * 10)), C.c_float(float(np.float32(dt))), C.c_float(tmax), C.byref(used))
    assert bad == 0
    assert used.value > 0.85 * n, "the closed form should apply to nearly every hop"


@pytest.mark.parametrize("dt", [1.73205080757 / 1024, 1.73205080757 / 4096, 2.0 ** -9, 1.5 * 2.0 ** -9, 1.73205080757 / 256])
@pytest.mark.parametrize("tmax", [0.9, 3.0, 30.0])
def test_cached_hop_sequences_equal_the_loop(shim, dt, tmax):
    """ngp_lattice_enter + ngp_lattice_hop_cached used the way the march kernel uses them: one ray, many samples and hops in a row"""
    used = C.c_long()
    n_rays, hops = 20000, 60
    bad = shim.run_cached(C.c_long(n_rays), C.c_int(hops), C.c_ulonglong(int(dt * 1e9) + int(tmax * 10) + 7), C.c_float(float(np.float32(dt))), C.c_float(tmax), C.byref(used))
    assert bad == 0
    assert used.value > 0.6 * n_rays * hops * 0.75


def test_hop_declines_what_it_cannot_prove(shim):
    out = C.c_float()
    dt = float(np.float32(1.73205080757 / 1024))
    assert shim.hop(0.999, 1.004, dt, C.byref(out)) == 0          # lands beyond the top of the binade [0.5, 1)
    assert shim.hop(0.0, 0.1, dt, C.byref(out)) == 0
    assert shim.hop(0.7, float("inf"), dt, C.byref(out)) == 0
    # a tie: dt exactly half an ulp of the binade above a multiple of it
    t = 1.0
    tie = float(np.float32(2.0 ** -10 + 2.0 ** -24))
    assert shim.hop(t, 1.01, tie, C.byref(out)) == 0
    # the do-while always takes one step, even when the target is behind
    assert shim.hop(0.7, 0.6, dt, C.byref(out)) == 1 and out.value == shim.loop(0.7, 0.6, dt)
