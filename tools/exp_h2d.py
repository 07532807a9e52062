#!/usr/bin/env python
"""Experiment: host->device bandwidth from pinned memory, alone and while the encoders keep the GPU busy."""
import os, sys, time, threading
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import govideocompressor_b200 as gvc
import bench
from concurrent.futures import ThreadPoolExecutor

gvc.init(0)
N = 8
seg = bench.synth_segment(bench.workload("cfg2", "testsrc"), 0)
host = [torch.from_numpy(seg).pin_memory() for _ in range(N)]
dev = [torch.empty_like(h, device="cuda") for h in host]
streams = [torch.cuda.Stream() for _ in range(N)]
nbytes = sum(h.numel() for h in host)

def copies(reps, nstreams):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for r in range(reps):
        for k in range(N):
            with torch.cuda.stream(streams[k % nstreams]):
                dev[k].copy_(host[k], non_blocking=True)
    torch.cuda.synchronize()
    return nbytes * reps / (time.perf_counter() - t0) / 1e9

copies(1, 1)
print(flush=True); print("H2D alone, 1 stream : %.1f GB/s" % copies(5, 1))
print("H2D alone, 8 streams: %.1f GB/s" % copies(5, 8))

encs = [gvc.Encoder(1920, 1080, 1280, 720, crf=23, host_threads=1) for _ in range(N)]
src = dev[0].clone()
stop = False
count = [0] * N
def spin(k):
    while not stop:
        encs[k].encode_device(src.data_ptr(), 38)
        count[k] += 1
ths = [threading.Thread(target=spin, args=(k,), daemon=True) for k in range(N)]
for t in ths: t.start()
def copies_nosync(reps, nstreams):
    # torch.cuda.synchronize() would wait for the encoder streams too: wait on the copy streams only
    t0 = time.perf_counter()
    for r in range(reps):
        for k in range(N):
            with torch.cuda.stream(streams[k % nstreams]):
                dev[k].copy_(host[k], non_blocking=True)
    for s_ in streams:
        s_.synchronize()
    return nbytes * reps / (time.perf_counter() - t0) / 1e9
try:
    time.sleep(0.5)
    c0 = sum(count); t0 = time.perf_counter()
    bw1 = copies_nosync(10, 1)
    c1 = sum(count); t1 = time.perf_counter()
    bw8 = copies_nosync(10, 8)
    c2 = sum(count); t2 = time.perf_counter()
    time.sleep(0.5)
    c3 = sum(count); t3 = time.perf_counter()
finally:
    stop = True
for t in ths: t.join(10)
print("H2D with encoders busy, 1 stream : %.1f GB/s while encoding %.0f frames/s" % (bw1, (c1 - c0) * 38 / (t1 - t0)))
print("H2D with encoders busy, 8 streams: %.1f GB/s while encoding %.0f frames/s" % (bw8, (c2 - c1) * 38 / (t2 - t1)))
print("encoders alone: %.0f frames/s" % ((c3 - c2) * 38 / (t3 - t2)))
