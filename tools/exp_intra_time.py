#!/usr/bin/env python
"""Experiment: CUDA-event time of the stages of one cfg2 segment alone on the GPU (stage pass of bench.py, stand-alone)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import govideocompressor_b200 as gvc
import bench
gvc.init(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
src = torch.from_numpy(bench.synth_segment(bench.workload("cfg2", "testsrc"), 0, n)).cuda()
enc = gvc.Encoder(1920, 1080, 1280, 720, crf=23)
for _ in range(2):
    enc.encode_device(src.data_ptr(), n); enc.sync()
enc.set_timing(True)
for _ in range(5):
    enc.encode_device(src.data_ptr(), n)
st = enc.stage_times()
for k, (ms, cnt) in st.items():
    if cnt: print("%-13s %4d launches  %8.1f us each" % (k, cnt, ms / cnt * 1e3))
