"""Scratch experiment: SM placement of the intra wavefront's CTAs, with and without a printf in the kernel."""
import subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
src = os.path.join(ROOT, "govideocompressor_b200", "csrc", "gvc_enc.cu")
orig = open(src).read()
END = '''    intra_row_done(lane, a, row_count);'''
SMID = '''    { unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid)); if (lane == 0) a.mail[(size_t)(a.mbh - 1) * a.mbw * 8 + my] = 0x100000000ull | smid; }
'''
RUN = r'''
import sys, os, ctypes, time
ROOT = %r
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import enc_util as E
from gpu_enc import Dev
dev = Dev(); lib = dev.lib; vp = ctypes.c_void_p
g = E.geom(1280, 720); W, H, nmb = g["w16"], g["h16"], g["mbw"] * g["mbh"]
ref_yuv, cur_yuv = E.make_pair(W, H, seed=5)
cur = E.Frame(g, cur_yuv, with_hpel=False)
cy, cu, cv = dev.up(cur.Y), dev.up(cur.U), dev.up(cur.V)
ry, ru, rv = dev.up(E.luma_plane(g)), dev.up(E.chroma_plane(g)), dev.up(E.chroma_plane(g))
co = torch.zeros(nmb * 26 * 16, dtype=torch.int16, device="cuda"); nz = torch.zeros(nmb, dtype=torch.int32, device="cuda")
modes = torch.zeros(nmb, dtype=torch.uint8, device="cuda"); scratch = torch.zeros(nmb * 8, dtype=torch.int64, device="cuda")
def run():
    lib.gvc_encode_intra(vp(cy.ptr), vp(cu.ptr), vp(cv.ptr), vp(ry.ptr), vp(ru.ptr), vp(rv.ptr), g["sy"], g["sc"], W, H, 20, vp(co.data_ptr()), vp(nz.data_ptr()), vp(modes.data_ptr()), vp(scratch.data_ptr()), None)
for _ in range(3): run()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(5): run()
b.record(); torch.cuda.synchronize()
sm = (scratch.cpu().numpy()[(g["mbh"] - 1) * g["mbw"] * 8:][:g["mbh"]] & 0xffffffff)
print("%%.1f us; smid per row: %%s" %% (a.elapsed_time(b) / 5 * 1e3, list(sm)))
''' % ROOT
variants = {
    "A-baseline+smid": lambda s: s.replace(END, SMID + END),
    "B-printf+smid": lambda s: s.replace(END, SMID + '    if (lane == 0 && my == 1000) printf("row %d\\n", my);\n' + END),
}
try:
    for name, fn in variants.items():
        s = fn(orig)
        if "#include <cstdio>" not in s:
            s = s.replace("#include <cuda_runtime.h>", "#include <cuda_runtime.h>\n#include <cstdio>", 1)
        open(src, "w").write(s)
        subprocess.check_call(["make", "-s", "-C", os.path.dirname(src)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        out = subprocess.run([sys.executable, "-c", RUN], capture_output=True, text=True)
        print(name, out.stdout.strip()[-900:], out.stderr.strip()[-300:], flush=True)
finally:
    open(src, "w").write(orig)
    subprocess.call(["make", "-s", "-C", os.path.dirname(src)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
