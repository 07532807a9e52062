"""Scratch experiment (not part of the product): which part of the timing instrumentation made intra_frame_kernel faster?"""
import subprocess, sys, os, shutil
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
src = os.path.join(ROOT, "govideocompressor_b200", "csrc", "gvc_enc.cu")
orig = open(src).read()
LOOP_HEAD = '''        intra_p0(lane, a, mx, my, sm);
        __syncwarp();'''
END = '''    intra_row_done(lane, a, row_count);'''
variants = {
    "A-baseline": lambda s: s,
    "B-printf-only": lambda s: s.replace(END, '    if (lane == 0 && my == 1000) printf("row %d\\n", my);\n' + END),
    "C-timer-start-end": lambda s: s.replace('    for (int mx = 0; mx < a.mbw; ++mx) {\n        for (int j = 0; j < 4; ++j) L.cur[j] = next[j];',
                                             '    unsigned long long t0, t1;\n    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));\n    for (int mx = 0; mx < a.mbw; ++mx) {\n        for (int j = 0; j < 4; ++j) L.cur[j] = next[j];')
                                    .replace(END, '    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));\n    if (t1 - t0 == 12345) a.modes[0] = 1;\n' + END),
    "D-timer-in-loop": lambda s: s.replace(LOOP_HEAD, LOOP_HEAD + '\n        if (mx == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));')
                                  .replace('    for (int mx = 0; mx < a.mbw; ++mx) {\n        for (int j = 0; j < 4; ++j) L.cur[j] = next[j];',
                                           '    unsigned long long tt = 0;\n    for (int mx = 0; mx < a.mbw; ++mx) {\n        for (int j = 0; j < 4; ++j) L.cur[j] = next[j];')
                                  .replace(END, '    if (lane == 0 && tt == 12345) printf("x %d\\n", my);\n' + END),
}
try:
    for name, fn in variants.items():
        s = fn(orig)
        if "#include <cstdio>" not in s:
            s = s.replace("#include <cuda_runtime.h>", "#include <cuda_runtime.h>\n#include <cstdio>", 1)
        open(src, "w").write(s)
        subprocess.check_call(["make", "-s", "-C", os.path.dirname(src)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "exp_intra_time.py")], capture_output=True, text=True).stdout
        print(name, [l for l in out.splitlines() if "intra" in l], flush=True)
finally:
    open(src, "w").write(orig)
    subprocess.call(["make", "-s", "-C", os.path.dirname(src)], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
