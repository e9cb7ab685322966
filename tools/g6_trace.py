#!/usr/bin/env python
"""Timeline of the decode step's GEMV launches (development): %globaltimer stamps per CTA of every gemv6 launch of one replayed
Llama-2-7B token (b200_debug_trace_begin / _read).  usage: python tools/g6_trace.py [model] [first_launch] [n_launches]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "ggml-hexagon_b200")]
import b200ops as B
import bench

model = sys.argv[1] if len(sys.argv) > 1 else "llama2-7b"
first = int(sys.argv[2]) if len(sys.argv) > 2 else 64
count = int(sys.argv[3]) if len(sys.argv) > 3 else 9
s = B.Session(0)
CAP = 400
B._ck(B.lib().b200_debug_trace_begin(s.h, CAP))
wl = bench.DecodeWorkload(B, s, model)
for _ in range(5):
    wl.step()
s.sync()
e0, e1 = s.event(), s.event()
s.record(e0)
for _ in range(10):
    wl.step()
s.record(e1)
print("step ms (traced build of the graph): %.4f" % (s.elapsed_ms(e0, e1) / 10))
buf = np.zeros(CAP * 4096, dtype=np.int64)
n = B.lib().b200_debug_trace_read(s.h, buf.ctypes.data_as(C.c_void_p), CAP)
tr = buf[:n * 4096].reshape(n, 1024, 4)
print("launches traced:", n)
base = None
names = ["start", "dep", "quant", "done"]
print("launch  ctas |  start min..max |    dep min..max |  quant min..max |   done min..max   (us, relative to the first start shown)")
for l in range(first, min(first + count, n)):
    t = tr[l]
    live = t[:, 0] > 0
    t = t[live].astype(np.float64) / 1e3
    if base is None:
        base = t[:, 0].min()
    t -= base
    print("%5d  %5d | %7.2f %7.2f | %7.2f %7.2f | %7.2f %7.2f | %7.2f %7.2f" % (l, live.sum(), t[:, 0].min(), t[:, 0].max(), t[:, 1].min(), t[:, 1].max(),
                                                                             t[:, 2].min(), t[:, 2].max(), t[:, 3].min(), t[:, 3].max()))
# whole step: time between consecutive launches' last "done"
done = np.array([tr[l][tr[l][:, 0] > 0][:, 3].max() for l in range(n)], dtype=np.float64) / 1e3
if n > 8:
    per = np.diff(done)
    print("done-to-done per launch (us), launches %d..%d:" % (first, first + count), np.round(per[first - 1:first + count - 1], 2))
    print("whole step from stamps: %.1f us over %d launches" % (done[-1] - tr[0][tr[0][:, 0] > 0][:, 0].min() / 1e3, n))
s.close()
