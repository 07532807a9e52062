#!/usr/bin/env python
"""Experiment: device-resident throughput of the path against the number of segments in flight.

    python tools/exp_streams.py 1 2 4 8 12 16                       # cfg2 shape: 1080p -> 720p, 38 pictures, me_range 16
    GVC_SHAPE=3840x2160:1920x1080:14:32 python tools/exp_streams.py 1 4 8    # cfg3 shape (src:dst:pictures:me_range)
"""
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np
import torch

import govideocompressor_b200 as gvc
from util_inputs import make_frame

SW, SH, DW, DH, N, MERANGE = 1920, 1080, 1280, 720, 38, 16
if os.environ.get("GVC_SHAPE"):
    a, b, c, d = os.environ["GVC_SHAPE"].split(":")
    (SW, SH), (DW, DH), N, MERANGE = map(int, a.split("x")), map(int, b.split("x")), int(c), int(d)


def main():
    counts = [int(a) for a in sys.argv[1:]] or [1, 2, 4, 8, 16]
    gvc.init(0)
    seg = np.empty((N, SW * SH * 3 // 2), np.uint8)
    for f in range(N):
        seg[f] = np.concatenate([p.reshape(-1) for p in make_frame(SW, SH, "yuv420p", "testsrc", 0, t=f)])
    src = torch.from_numpy(seg).cuda()
    steps = 6
    for s in counts:
        encs = [gvc.Encoder(SW, SH, DW, DH, crf=23, merange=MERANGE, host_threads=1) for _ in range(s)]
        pool = ThreadPoolExecutor(s)
        run = lambda n: list(pool.map(lambda k: [encs[k].encode_device(src.data_ptr(), N) for _ in range(n)], range(s)))
        run(2)
        for e in encs:
            e.sync()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        run(steps)
        for e in encs:
            e.sync()
        dt = time.perf_counter() - t0
        print("segments in flight %2d: %8.0f frames/s  (%.2f ms per segment-step)" % (s, s * N * steps / dt, dt / steps * 1e3), flush=True)
        for e in encs:
            e.close()
        pool.shutdown()


if __name__ == "__main__":
    main()
