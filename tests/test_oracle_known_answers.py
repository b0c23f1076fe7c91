"""CPU: pins the oracle with analytic known answers (the reference has no golden vectors)."""
import ctypes as C

import numpy as np

from util import pcg_floats


def test_pcg32_matches_python_restatement(orc):
    r = orc.Pcg32(1337)
    ref = pcg_floats(8, 1337)
    got = np.array([r.next_float() for _ in range(8)], np.float32)
    assert np.array_equal(ref, got)


def test_pcg32_advance_equals_stepping(orc):
    a, b = orc.Pcg32(43), orc.Pcg32(43)
    for _ in range(1000):
        a.next_uint()
    b.advance(1000)
    assert a.state == b.state
    b.advance(-1000 & ((1 << 63) - 1) if False else 0)  # advance(0) is the identity
    assert a.state == b.state


def test_trainer_seed_is_std_seed_seq(orc):
    # std::seed_seq{43}.generate -> pcg32{first}; value fixed by the C++ standard's algorithm
    r = orc.trainer_rng(43)
    assert r.inc == 3
    s2 = orc.trainer_rng(43)
    assert r.state == s2.state and r.state != orc.trainer_rng(44).state


def test_grid_levels_match_survey_constants(orc):
    L = orc.GridLevels(16, 19, 16)
    assert list(L.resolutions) == [16, 23, 31, 43, 59, 81, 112, 154, 213, 295, 407, 562, 777, 1073, 1483, 2048]
    assert L.n_entries == 6098120 and L.n_params == 12196240
    sizes = np.diff(L.offsets.astype(np.int64))
    assert list(sizes[:5]) == [4096, 12168, 29792, 79512, 205384] and all(sizes[5:] == 1 << 19)
    assert abs(float(L.per_level_scale) - 1.381912879967776) < 1e-6
    big = orc.GridLevels(16, 22, 16)
    assert big.n_entries == 39474024


def test_hash_index_known_answers(orc):
    lib = orc.lib()
    T = 1 << 19
    # hashed level: fast_hash(1,1,1) = 1 ^ 19349663 ^ 83492791, then mod T
    assert lib.orc_grid_index(T, 2048, 1, 1, 1) == (1 ^ 19349663 ^ 83492791) % T
    assert lib.orc_grid_index(T, 2048, 5, 0, 0) == 5
    assert lib.orc_grid_index(T, 2048, 0, 3, 0) == (3 * 19349663) % (1 << 32) % T
    # dense level: x + y*res + z*res^2, wrapped by the padded size
    assert lib.orc_grid_index(4096, 16, 3, 2, 1) == 3 + 2 * 16 + 1 * 256
    assert lib.orc_grid_index(4096, 16, 16, 15, 15) == (16 + 15 * 16 + 15 * 256) % 4096
    # grid.h:88-95 with a uint32_t stride: at T = 2^22 the finest level has res^2 == size, the next product is 2^33 == 0 (mod 2^32),
    # "size < stride" is false and the level is indexed densely -- the z term vanishes modulo the size
    T22 = 1 << 22
    assert lib.orc_grid_index(T22, 2048, 7, 9, 1234) == (7 + 9 * 2048) % T22
    assert lib.orc_grid_index(T22, 1483, 7, 9, 12) == (7 ^ (9 * 19349663) % (1 << 32) ^ (12 * 83492791) % (1 << 32)) % T22


def test_hash_encode_constant_table_is_partition_of_unity(orc):
    # a table holding the same value everywhere must interpolate to that value (weights sum to 1) up to fp16 rounding
    L = orc.GridLevels(16, 19, 16)
    table = np.full(L.n_params, np.float16(0.5)).view(np.uint16)
    xyz = pcg_floats(3 * 64, 7).reshape(64, 3).copy()
    out, idx = orc.hash_encode_fwd(xyz, 3, 64, table, L, want_indices=True)
    assert np.all(np.abs(out.astype(np.float32) - 0.5) <= 4 * 2.0 ** -11)
    sizes = np.diff(L.offsets.astype(np.int64))
    assert np.all(idx < sizes[None, :, None])


def test_hash_encode_bwd_is_adjoint_of_fwd(orc):
    # <dy, fwd(table)> == <bwd(dy), table> for a linear operator (up to fp16 rounding of terms)
    L = orc.GridLevels(4, 12, 4)
    n = 256
    g = np.random.Generator(np.random.PCG64(3))
    table = (g.random(L.n_params, dtype=np.float32) - 0.5).astype(np.float16)
    xyz = g.random((n, 3), dtype=np.float32)
    dy = (g.random((n, 2 * L.n_levels), dtype=np.float32) - 0.5).astype(np.float16)
    y = orc.hash_encode_fwd(xyz, 3, n, table.view(np.uint16), L).astype(np.float64)
    grad = orc.hash_encode_bwd(xyz, 3, n, dy, L).astype(np.float64)
    lhs = float((dy.astype(np.float64) * y).sum())
    rhs = float((grad * table.astype(np.float64)).sum())
    assert abs(lhs - rhs) < 2e-2 * max(1.0, abs(lhs))


def test_morton_and_sh_axes(orc):
    sh = orc.sh4(np.array([[1.0, 0.5, 0.5], [0.5, 1.0, 0.5], [0.5, 0.5, 1.0]], np.float32)).astype(np.float32)
    # directions +x, +y, +z: l=1 band is (-c*y, c*z, -c*x) with c = sqrt(3/(4 pi))
    c = 0.4886025
    assert np.allclose(sh[:, 0], 0.2820948, atol=1e-3)
    assert np.allclose(sh[0, 1:4], [0, 0, -c], atol=1e-3)
    assert np.allclose(sh[1, 1:4], [-c, 0, 0], atol=1e-3)
    assert np.allclose(sh[2, 1:4], [0, c, 0], atol=1e-3)
    assert np.allclose(sh[2, 6], 0.9461747 - 0.3153916, atol=1e-3)   # z^2 = 1


def test_ld_random_val_range_and_determinism(orc):
    lib = orc.lib()
    v = [lib.orc_ld_random_val(C.c_uint32(i), C.c_uint32(0xDEADBEEF), C.c_uint32(0)) for i in range(64)]
    assert all(0.0 <= x < 1.0 for x in v) and len(set(v)) > 60
    assert v[3] == lib.orc_ld_random_val(C.c_uint32(3), C.c_uint32(0xDEADBEEF), C.c_uint32(0))
