"""test helper: a minimal PNG writer with full control over what the reference's stb_image decoder has to handle -- colour type,
bit depth, the filter type of every row, Adam7 interlacing, PLTE / tRNS chunks -- plus the RGBA8 image stb_image's
stbi_load(path, ..., 4) yields for it (dependencies/stb_image/stb_image.h conversions, restated here as the expectation)."""
import struct
import zlib

import numpy as np

CHANNELS = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}
ADAM7 = [(0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)]


def _chunk(tag, body):
    return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body) & 0xFFFFFFFF)


def _pack_rows(samples, depth):
    """samples [h][w*ch] integer -> list of byte rows"""
    rows = []
    for r in samples:
        if depth == 8:
            rows.append(np.asarray(r, np.uint8).tobytes())
        elif depth == 16:
            rows.append(np.asarray(r, ">u2").tobytes())
        else:
            bits = np.zeros(((len(r) * depth + 7) // 8) * 8, np.uint8)
            for k, v in enumerate(r):
                for b in range(depth):
                    bits[k * depth + b] = (int(v) >> (depth - 1 - b)) & 1
            rows.append(np.packbits(bits).tobytes())
    return rows


def _filter_rows(rows, bpp, filters):
    out, prev = b"", None
    for y, row in enumerate(rows):
        f = filters[y % len(filters)]
        cur = np.frombuffer(row, np.uint8).astype(np.int32)
        up = np.frombuffer(prev, np.uint8).astype(np.int32) if prev is not None else np.zeros_like(cur)
        left = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]]) if len(cur) > bpp else np.zeros_like(cur)
        ul = np.concatenate([np.zeros(bpp, np.int32), up[:-bpp]]) if len(cur) > bpp else np.zeros_like(cur)
        if f == 0:
            pred = 0
        elif f == 1:
            pred = left
        elif f == 2:
            pred = up
        elif f == 3:
            pred = (left + up) >> 1
        else:
            p = left + up - ul
            pa, pb, pc = np.abs(p - left), np.abs(p - up), np.abs(p - ul)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, up, ul))
        out += bytes([f]) + ((cur - pred) & 255).astype(np.uint8).tobytes()
        prev = row
    return out


def write_png(path, samples, ctype, depth=8, filters=(0,), interlace=False, palette=None, trns=None):
    """samples: integer array [h][w][channels] (channels = CHANNELS[ctype]) of raw sample values at `depth` bits"""
    samples = np.asarray(samples)
    h, w, ch = samples.shape
    assert ch == CHANNELS[ctype]
    bpp = max(1, ch * depth // 8)
    raw = b""
    passes = ADAM7 if interlace else [(0, 0, 1, 1)]
    for x0, y0, dx, dy in passes:
        sub = samples[y0::dy, x0::dx]
        if sub.shape[0] == 0 or sub.shape[1] == 0:
            continue
        raw += _filter_rows(_pack_rows(sub.reshape(sub.shape[0], -1), depth), bpp, filters)
    data = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 1 if interlace else 0))
    if palette is not None:
        data += _chunk(b"PLTE", np.asarray(palette, np.uint8).tobytes())
    if trns is not None:
        data += _chunk(b"tRNS", bytes(trns))
    comp = zlib.compress(raw, 6)
    half = len(comp) // 2                       # two IDAT chunks: the decoder has to concatenate them
    data += _chunk(b"IDAT", comp[:half]) + _chunk(b"IDAT", comp[half:]) + _chunk(b"IEND", b"")
    open(path, "wb").write(data)


def expected_rgba8(samples, ctype, depth=8, palette=None, trns=None):
    """what stbi_load(..., req_comp = 4) returns: 16-bit -> high byte, sub-byte gray scaled to 0..255, gray -> (g, g, g), alpha 255
    unless given; tRNS = colour key (gray / RGB: alpha 0 where the pixel matches) or per-entry palette alphas"""
    s = np.asarray(samples).astype(np.int64)
    h, w, _ = s.shape
    scale = {1: 0xFF, 2: 0x55, 4: 0x11, 8: 1}.get(depth, 1)
    to8 = (lambda v: v >> 8) if depth == 16 else (lambda v: v * (scale if ctype == 0 else 1))
    out = np.zeros((h, w, 4), np.uint8)
    if ctype == 3:
        pal = np.zeros((256, 4), np.uint8)
        pal[:, 3] = 255
        p = np.asarray(palette, np.uint8).reshape(-1, 3)
        pal[:len(p), :3] = p
        if trns is not None:
            pal[:len(trns), 3] = list(trns)
        return pal[s[..., 0]]
    if ctype in (0, 4):
        out[..., 0] = out[..., 1] = out[..., 2] = to8(s[..., 0])
        out[..., 3] = to8(s[..., 1]) if ctype == 4 else 255
        if ctype == 0 and trns is not None:
            key = (trns[0] << 8) | trns[1]
            match = (s[..., 0] == key) if depth == 16 else (to8(s[..., 0]) == ((key & 255) * (scale if depth < 8 else 1)) & 255)
            out[..., 3] = np.where(match, 0, 255)
        return out
    out[..., :3] = to8(s[..., :3])
    out[..., 3] = to8(s[..., 3]) if ctype == 6 else 255
    if ctype == 2 and trns is not None:
        key = [(trns[2 * k] << 8) | trns[2 * k + 1] for k in range(3)]
        match = np.all(s[..., :3] == key, axis=-1) if depth == 16 else np.all(out[..., :3] == [k & 255 for k in key], axis=-1)
        out[..., 3] = np.where(match, 0, 255)
    return out


def cases(seed=3, w=23, h=17):
    """(name, kwargs for write_png) covering what the decoder distinguishes"""
    g = np.random.Generator(np.random.PCG64(seed))
    r8 = lambda ch: g.integers(0, 256, (h, w, ch))
    smooth = np.clip(np.add.outer(np.arange(h) * 9, np.arange(w) * 7)[..., None] + g.integers(0, 12, (h, w, 4)), 0, 255)    # gradients: the filters matter
    out = [("rgba8_all_filters", dict(samples=smooth, ctype=6, filters=(0, 1, 2, 3, 4)))]
    for f in range(5):
        out.append((f"rgba8_filter{f}", dict(samples=r8(4), ctype=6, filters=(f,))))
    out += [("rgb8", dict(samples=r8(3), ctype=2, filters=(4, 1))), ("gray8", dict(samples=r8(1), ctype=0, filters=(3,))),
            ("graya8", dict(samples=r8(2), ctype=4, filters=(2, 4))),
            ("rgba16", dict(samples=g.integers(0, 65536, (h, w, 4)), ctype=6, depth=16, filters=(1, 4, 3))),
            ("rgb16_key", dict(samples=np.where(g.random((h, w, 1)) < 0.3, [[[300, 20000, 65535]]], g.integers(0, 65536, (h, w, 3))), ctype=2, depth=16, trns=[1, 44, 78, 32, 255, 255])),
            ("rgb8_key", dict(samples=np.where(g.random((h, w, 1)) < 0.3, [[[10, 200, 30]]], r8(3)), ctype=2, trns=[0, 10, 0, 200, 0, 30])),
            ("gray1", dict(samples=g.integers(0, 2, (h, w, 1)), ctype=0, depth=1)), ("gray2", dict(samples=g.integers(0, 4, (h, w, 1)), ctype=0, depth=2, filters=(0, 2))),
            ("gray4_key", dict(samples=g.integers(0, 16, (h, w, 1)), ctype=0, depth=4, trns=[0, 5])), ("gray16", dict(samples=g.integers(0, 65536, (h, w, 1)), ctype=0, depth=16, filters=(4,))),
            ("palette8_alpha", dict(samples=g.integers(0, 40, (h, w, 1)), ctype=3, palette=g.integers(0, 256, (40, 3)), trns=list(g.integers(0, 256, 25)))),
            ("palette4", dict(samples=g.integers(0, 16, (h, w, 1)), ctype=3, depth=4, palette=g.integers(0, 256, (16, 3)))),
            ("palette2", dict(samples=g.integers(0, 4, (h, w, 1)), ctype=3, depth=2, palette=g.integers(0, 256, (4, 3)))),
            ("rgba8_adam7", dict(samples=smooth, ctype=6, filters=(4, 3, 1), interlace=True)), ("gray1_adam7", dict(samples=g.integers(0, 2, (h, w, 1)), ctype=0, depth=1, interlace=True)),
            ("rgba16_adam7", dict(samples=g.integers(0, 65536, (h, w, 4)), ctype=6, depth=16, filters=(2,), interlace=True))]
    return out
