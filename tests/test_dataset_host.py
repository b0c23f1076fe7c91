"""CPU tests of the host side of the dataset / evaluation path: the PNG decoder behind load_training_data (the reference decodes
with stb_image, internal/dataset/nerf_synthetic.cu:82) against hand-built PNGs of every colour type, bit depth, filter type and
interlacing, and against PIL's decoder; the image metrics of the evaluation loop (experiment/nerf_synthetic/python/common.py)."""
import importlib
import os

import numpy as np
import pytest

import pngutil

PKG = "instant-ngp-nerf-minimal_b200"


@pytest.fixture(scope="module")
def pyngp():
    pkg = importlib.import_module(PKG)
    return pkg.load_pyngp()                 # host code only is exercised here; the module loads without a GPU


@pytest.mark.parametrize("name,kw", pngutil.cases(), ids=[c[0] for c in pngutil.cases()])
def test_png_decoder_matches_the_stb_image_conversion(pyngp, tmp_path, name, kw):
    path = os.path.join(tmp_path, name + ".png")
    pngutil.write_png(path, **kw)
    got = np.array(pyngp.decode_png(path))
    want = pngutil.expected_rgba8(kw["samples"], kw["ctype"], kw.get("depth", 8), kw.get("palette"), kw.get("trns"))
    assert got.shape == want.shape and got.dtype == np.uint8
    assert np.array_equal(got, want)
    if kw.get("depth", 8) == 8 and kw.get("trns") is None:
        import PIL.Image
        assert np.array_equal(got, np.asarray(PIL.Image.open(path).convert("RGBA"))), "an independent decoder agrees on the 8-bit cases"


def test_png_decoder_reads_what_a_real_encoder_writes(pyngp, tmp_path):
    import PIL.Image
    g = np.random.Generator(np.random.PCG64(5))
    img = np.clip(np.add.outer(np.arange(64) * 3, np.arange(80) * 2)[..., None] + g.integers(0, 30, (64, 80, 4)), 0, 255).astype(np.uint8)
    for mode, arr in (("RGBA", img), ("RGB", img[..., :3]), ("L", img[..., 0]), ("LA", img[..., :2])):
        path = os.path.join(tmp_path, mode + ".png")
        PIL.Image.fromarray(arr, mode).save(path, optimize=True)
        assert np.array_equal(np.array(pyngp.decode_png(path)), np.asarray(PIL.Image.open(path).convert("RGBA"))), mode


def test_png_decoder_rejects_garbage(pyngp, tmp_path):
    path = os.path.join(tmp_path, "bad.png")
    open(path, "wb").write(b"\x89PNG\r\n\x1a\n" + b"\0" * 40)
    with pytest.raises(RuntimeError):
        pyngp.decode_png(path)
    open(path, "wb").write(b"not a png at all")
    with pytest.raises(RuntimeError):
        pyngp.decode_png(path)
    pngutil.write_png(path, samples=np.zeros((4, 4, 4), np.int64), ctype=6)
    data = open(path, "rb").read()
    open(path, "wb").write(data[:len(data) // 2])
    with pytest.raises(RuntimeError):
        pyngp.decode_png(path)


def test_eval_metrics_follow_common_py():
    run = importlib.import_module(PKG + ".run")
    from scipy.ndimage import convolve1d
    g = np.random.Generator(np.random.PCG64(1))
    a, b = g.random((40, 56, 3)), g.random((40, 56, 3))
    k = np.array([0.120078, 0.233881, 0.292082, 0.233881, 0.120078])
    for axis in (0, 1):
        assert np.allclose(run._convolve1d_reflect(a[..., 0], k, axis), convolve1d(a[..., 0], k, axis=axis), atol=1e-12)

    def ref_ssim(a, b):                       # common.py:134-151, with scipy as the reference imports it
        blur = lambda x: convolve1d(convolve1d(x, k, axis=0), k, axis=1)
        a, b = run.luminance(a), run.luminance(b)
        mA, mB = blur(a), blur(b)
        sA, sB, sAB = blur(a * a) - mA ** 2, blur(b * b) - mB ** 2, blur(a * b) - mA * mB
        return (2.0 * mA * mB + 1e-4) / (mA * mA + mB * mB + 1e-4) * (2.0 * sAB + 9e-4) / (sA + sB + 9e-4)
    assert np.allclose(run.ssim_map(a, b), ref_ssim(a, b), atol=1e-12)
    assert run.compute_error("SSIM", a, a) == pytest.approx(1.0)
    assert run.compute_error("MSE", a, b) == pytest.approx(float(np.mean((a - b) ** 2)))
    assert run.mse2psnr(0.01) == pytest.approx(20.0)
    x = np.linspace(0, 1, 1001)
    assert np.allclose(run.srgb_to_linear(run.linear_to_srgb(x)), x, atol=1e-6)
    assert run.linear_to_srgb(np.array(0.0031308)) == pytest.approx(12.92 * 0.0031308, abs=1e-6)
