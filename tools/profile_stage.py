#!/usr/bin/env python
"""Single-threaded driver for ncu: runs one stage of the path a few times with plain launches (no CUDA graph, no host
threads), so `ncu -k regex:<kernel>` sees exactly the kernels of that stage.

    python tools/profile_stage.py resize [n_frames]      # resize_fused_kernel, 1080p -> 720p
    python tools/profile_stage.py segment [n_frames]     # every encoder kernel of one cfg2 segment
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np
import torch

import govideocompressor_b200 as gvc
from util_inputs import make_frame

SW, SH, DW, DH = 1920, 1080, 1280, 720


def frames(n):
    seg = np.empty((n, SW * SH * 3 // 2), np.uint8)
    for f in range(n):
        seg[f] = np.concatenate([p.reshape(-1) for p in make_frame(SW, SH, "yuv420p", "testsrc", 0, t=f)])
    return seg


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "resize"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 38
    gvc.init(0)
    src = torch.from_numpy(frames(n)).cuda()
    if what == "resize":
        plan = gvc.ResizePlan(SW, SH, DW, DH)
        dst = torch.empty((n, plan.frame_bytes(DW, DH)), dtype=torch.uint8, device="cuda")
        for _ in range(4):
            plan.resize_device(src, dst, n)
        torch.cuda.synchronize()
    else:
        enc = gvc.Encoder(SW, SH, DW, DH, crf=23, disable_graph=True)
        for _ in range(3):
            enc.encode_device(src.data_ptr(), n)
            enc.sync()
        enc.close()
    print("ok")


if __name__ == "__main__":
    main()
