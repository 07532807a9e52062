"""Synthetic I420 sequences for the encoder tests."""
import numpy as np

import enc_util as E


def moving_sequence(w, h, n, seed=1, noise=0, static_left=False, fast=False):
    """Textured background translating (2,1) px/frame (or (7,-5) with fast=True), an independently moving 32x32 patch,
    optional noise, optionally a frozen left third (exercises P_Skip)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    m = 96
    by = E.smooth_field(h + 2 * m, w + 2 * m, seed, 8)
    bu = E.smooth_field(h // 2 + m, w // 2 + m, seed + 1, 4)
    bv = E.smooth_field(h // 2 + m, w // 2 + m, seed + 2, 4)
    patch = E.smooth_field(48, 48, seed + 3, 2)[:32, :32]
    cw, ch = (w + 1) // 2, (h + 1) // 2
    frames = np.zeros((n, w * h + 2 * cw * ch), np.uint8)
    sx, sy = (7, -5) if fast else (2, 1)
    first = None
    for f in range(n):
        ox, oy = m // 2 + sx * f, m // 2 + sy * f
        y = by[oy:oy + h, ox:ox + w].copy()
        u = bu[oy // 2:oy // 2 + ch, ox // 2:ox // 2 + cw].copy()
        v = bv[oy // 2:oy // 2 + ch, ox // 2:ox // 2 + cw].copy()
        px, py = (w // 2 + 3 * f) % max(1, w - 32), (h // 3 + 2 * f) % max(1, h - 32)
        y[py:py + 32, px:px + 32] = patch
        if noise:
            y = np.clip(y.astype(int) + rng.integers(-noise, noise + 1, y.shape), 0, 255).astype(np.uint8)
        if static_left:
            if first is None:
                first = (y.copy(), u.copy(), v.copy())
            y[:, : w // 3] = first[0][:, : w // 3]
            u[:, : cw // 3] = first[1][:, : cw // 3]
            v[:, : cw // 3] = first[2][:, : cw // 3]
        frames[f] = np.concatenate([y.reshape(-1), u.reshape(-1), v.reshape(-1)])
    return frames


def split_i420(frame, w, h):
    cw, ch = (w + 1) // 2, (h + 1) // 2
    return frame[: w * h].reshape(h, w), frame[w * h: w * h + cw * ch].reshape(ch, cw), frame[w * h + cw * ch:].reshape(ch, cw)


def psnr(a, b):
    mse = np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)
    return 99.0 if mse == 0 else 10 * np.log10(255.0 ** 2 / mse)


def parameter_sets(annexb):
    """(SPS bytes, PPS bytes) of an Annex B stream: the first NAL units of type 7 and 8, start codes stripped."""
    b, found = bytes(annexb), {}
    i = b.find(b"\x00\x00\x01")
    while i >= 0:
        j = b.find(b"\x00\x00\x01", i + 3)
        nal = b[i + 3: (j - 1 if j > 0 and b[j - 1] == 0 else j) if j >= 0 else len(b)]
        if nal and (nal[0] & 31) in (7, 8):
            found.setdefault(nal[0] & 31, nal)
        i = j
    return found.get(7), found.get(8)


def write_concat_list(directory, names):
    """filelist.txt as server.go:325-347 writes it (`file 'N.mp4'` lines).  The operator's concat.sh (server.go:357) names the
    demuxer with `-f concat`; a library caller gets the same demuxer by probing, which needs the format's header line."""
    path = directory + "/filelist.txt"
    with open(path, "w") as f:
        f.write("ffconcat version 1.0\n")
        for n in names:
            f.write("file '%s'\n" % n)
    return path
