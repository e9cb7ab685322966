#!/usr/bin/env python
"""Summarise `nvcc -Xptxas -v` logs: registers / spills / barriers per kernel (run after make)."""
import re, subprocess, sys, glob
for path in sys.argv[1:] or glob.glob('ggml-hexagon_b200/lib/*.ptxas.log'):
    log = open(path).read()
    rows = []
    for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'\n.*\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\nptxas info\s+: Used (\d+) registers(?:, used (\d+) barriers)?(?:, (\d+) bytes smem)?", log):
        name = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip()
        name = re.sub(r'\(.*', '', name).replace('void b200::', '')
        rows.append((int(m.group(5)), name, int(m.group(3)), m.group(7) or '0'))
    print(path, len(rows), 'kernels')
    for r in sorted(rows, reverse=True):
        print('  regs=%3d spill=%d smem=%s %s' % (r[0], r[2], r[3], r[1]))
