"""CPU: the C-ABI library loads and exports every symbol include/ngp_b200.h declares; host-only entry points agree with the oracle."""
import ctypes as C
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "ngp_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ngp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(capi):
    lib = capi.load()
    names = declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in ngp_b200.h but not exported"
        assert n in capi.SIGNATURES, f"{n} has no ctypes signature"
    assert lib.ngp_abi_version() == 1


def test_grid_desc_matches_oracle(capi, orc):
    for (nl, log2t, base) in [(16, 19, 16), (16, 22, 16), (8, 15, 4)]:
        d, n_entries = capi.make_grid_desc(nl, log2t, base)
        L = orc.GridLevels(nl, log2t, base)
        assert n_entries == L.n_entries
        assert list(d.offsets[: nl + 1]) == list(L.offsets)
        assert list(d.resolutions[:nl]) == list(L.resolutions)
        assert np.array_equal(np.array(d.scales[:nl], np.float32), L.scales)


def test_errors_are_reported_not_swallowed(capi):
    lib = capi.load()
    d, _ = capi.make_grid_desc()
    rc = lib.ngp_hash_encode_fwd(None, None, 3, 16, None, C.byref(d), None, 32, None)
    assert rc != 0 and b"null pointer" in lib.ngp_last_error()
    bad = capi.GridDesc()
    assert lib.ngp_grid_desc_init(C.byref(bad), 40, 19, 16, 1.38) == 0 and b"n_levels" in lib.ngp_last_error()


def test_sampler_cfg_constants(capi):
    c = capi.make_sampler_cfg(1)
    assert c.n_cascades == 2 and abs(c.min_step - 1.73205080757 / 1024) < 1e-9 and abs(c.max_step - 1.73205080757 * 2 / 128) < 1e-8
    assert capi.make_sampler_cfg(16).n_cascades == 5
