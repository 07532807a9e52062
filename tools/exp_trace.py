"""Developer aid: e2e timeline of the 8-slot host path with GVC_TRACE=1 (prints the library's trace to stderr)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("GVC_TRACE", "1")
import numpy as np, torch
import govideocompressor_b200 as gvc
import bench
from concurrent.futures import ThreadPoolExecutor
gvc.init(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
encs = [gvc.Encoder(1920, 1080, 1280, 720, crf=23, host_threads=int(os.environ.get("GVC_HOST_THREADS", "2"))) for _ in range(n)]
seg = bench.synth_segment(bench.workload("cfg2", "testsrc"), 0)
host = [torch.from_numpy(seg).pin_memory() for _ in range(n)]
pool = ThreadPoolExecutor(n)
def slot(k, steps):
    e, src = encs[k], host[k].data_ptr()
    e.upload_host(src, 38)
    for i in range(steps):
        e.encode_uploaded()
        if i + 1 < steps: e.upload_host(src, 38)
        e.finish_discard()
list(pool.map(lambda k: slot(k, 2), range(n)))
torch.cuda.synchronize()
print("---- timed", file=sys.stderr)
t0 = time.perf_counter()
STEPS = int(os.environ.get("GVC_STEPS", "4"))
list(pool.map(lambda k: slot(k, STEPS), range(n)))
dt = time.perf_counter() - t0
print("slots", n, "fps", n * 38 * STEPS / dt, file=sys.stderr)
