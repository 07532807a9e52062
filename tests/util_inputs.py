"""Deterministic synthetic inputs shared by the golden generator, the tests and bench.py."""
import numpy as np


def _chroma_dims(w, h):
    return (w + 1) // 2, (h + 1) // 2


def testsrc_planes(w, h, t=0, seed=0):
    """SURVEY.md §8(d) synthetic frame: colour bars + translating diagonal gradient (+2,+1 px/frame)
    + orbiting 64x64 box + one 128x128 LCG-noise patch.  Integer arithmetic only."""
    yy, xx = np.mgrid[0:h, 0:w].astype(np.int64)
    bars_y = np.array([235, 210, 170, 145, 106, 81, 41, 16], np.int64)
    bars_u = np.array([128, 16, 166, 54, 202, 90, 240, 128], np.int64)
    bars_v = np.array([128, 146, 16, 34, 222, 240, 110, 128], np.int64)
    bar = np.minimum((xx * 8) // max(w, 1), 7)
    Y = bars_y[bar]
    grad = ((xx - 2 * t) + (yy - t)) & 255
    Y = np.where(yy >= h // 2, (Y + grad) >> 1, Y)
    bx = (w // 2 - 32) + ((w // 4) * np.cos(2 * np.pi * t / 60.0)).astype(np.int64)
    by = (h // 2 - 32) + ((h // 4) * np.sin(2 * np.pi * t / 60.0)).astype(np.int64)
    box = (xx >= bx) & (xx < bx + 64) & (yy >= by) & (yy < by + 64)
    Y = np.where(box, 255 - Y, Y)
    px, py = w // 8, h // 8
    lcg = (((xx - px) + 131 * (yy - py) + 7919 * (seed + 1) + 31 * t) * 1103515245 + 12345) >> 16 & 255
    patch = (xx >= px) & (xx < px + 128) & (yy >= py) & (yy < py + 128)
    Y = np.where(patch, lcg, Y).astype(np.uint8)
    cw, ch = _chroma_dims(w, h)
    cbar = np.minimum((np.arange(cw, dtype=np.int64) * 8) // max(cw, 1), 7)
    U = np.broadcast_to(bars_u[cbar], (ch, cw)).astype(np.uint8).copy()
    V = np.broadcast_to(bars_v[cbar], (ch, cw)).astype(np.uint8).copy()
    return [Y, U, V]


def make_frame(w, h, fmt="yuv420p", kind="random", seed=0, t=0):
    """Returns planes: [Y,U,V] for yuv420p, [Y,UV] for nv12."""
    cw, ch = _chroma_dims(w, h)
    if kind == "random":
        rng = np.random.Generator(np.random.PCG64(seed))
        p = [rng.integers(0, 256, (h, w), dtype=np.uint8), rng.integers(0, 256, (ch, cw), dtype=np.uint8),
             rng.integers(0, 256, (ch, cw), dtype=np.uint8)]
    elif kind == "zeros":
        p = [np.zeros((h, w), np.uint8), np.zeros((ch, cw), np.uint8), np.zeros((ch, cw), np.uint8)]
    elif kind == "full":
        p = [np.full((h, w), 255, np.uint8), np.full((ch, cw), 255, np.uint8), np.full((ch, cw), 255, np.uint8)]
    elif kind == "checker":
        yy, xx = np.mgrid[0:h, 0:w]
        cy, cx = np.mgrid[0:ch, 0:cw]
        p = [(((xx + yy) & 1) * 255).astype(np.uint8), (((cx + cy) & 1) * 255).astype(np.uint8),
             ((((cx >> 1) + cy) & 1) * 255).astype(np.uint8)]
    elif kind == "testsrc":
        p = testsrc_planes(w, h, t, seed)
    elif kind == "pan":
        p = pan_planes(w, h, t, seed)
    else:
        raise ValueError(kind)
    if fmt == "nv12":
        uv = np.empty((ch, 2 * cw), np.uint8)
        uv[:, 0::2] = p[1]
        uv[:, 1::2] = p[2]
        return [p[0], uv]
    return p


_PAN_CACHE = {}


def pan_planes(w, h, t=0, seed=0):
    """Second content point for the bench: a band-limited texture (noise blown up 8x and box filtered, plus fine grain) that
    translates as a whole by (+3, +2) luma samples per picture -- global motion, so no macroblock is predicted by the zero
    vector and none leaves the motion search early; chroma moves with it."""
    key = (w, h, seed)
    if key not in _PAN_CACHE:
        _PAN_CACHE.clear()
        rng = np.random.Generator(np.random.PCG64(1000 + seed))
        m = 512                                    # room for the translation

        def field(hh, ww, scale, grain):
            small = rng.integers(0, 256, ((hh + m) // scale + 3, (ww + m) // scale + 3)).astype(np.float64)
            big = np.kron(small, np.ones((scale, scale)))
            k = scale
            c = np.cumsum(np.cumsum(np.pad(big, ((k, 0), (k, 0))), 0), 1)
            box = (c[k:, k:] - c[:-k, k:] - c[k:, :-k] + c[:-k, :-k]) / (k * k)
            return np.clip(box + rng.integers(-grain, grain + 1, box.shape), 0, 255).astype(np.uint8)
        cw, ch = _chroma_dims(w, h)
        _PAN_CACHE[key] = (field(h, w, 8, 10), field(ch, cw, 4, 3), field(ch, cw, 4, 3))
    Y, U, V = _PAN_CACHE[key]
    cw, ch = _chroma_dims(w, h)
    ox, oy = (3 * t) % 500, (2 * t) % 500
    ox, oy = ox - (ox & 1), oy - (oy & 1)          # even: chroma moves by whole samples too
    return [Y[oy:oy + h, ox:ox + w].copy(), U[oy // 2:oy // 2 + ch, ox // 2:ox // 2 + cw].copy(), V[oy // 2:oy // 2 + ch, ox // 2:ox // 2 + cw].copy()]
