#!/usr/bin/env python
"""Reads a B200_STREAM_TRACE dump (decode_stream.cu development trace) and prints, per stage position in a layer,
the mean SM-clock cycles thread 0 spent: waiting for inputs, working, publishing."""
import sys
import numpy as np

raw = open(sys.argv[1], "rb").read()
G, S = np.frombuffer(raw[:8], dtype=np.int32)
t = np.frombuffer(raw[8:], dtype=np.int64).reshape(G, S, 4).astype(np.float64)
period = int(sys.argv[2]) if len(sys.argv) > 2 else 4
wait = t[:, :, 1] - t[:, :, 0]
work = t[:, :, 2] - t[:, :, 1]
pub = t[:, :, 3] - t[:, :, 2]
tot = t[:, 1:, 0] - t[:, :-1, 0]
print("grid %d, stages %d; cycles (mean over CTAs and layers); per stage slot in a period of %d" % (G, S, period))
for p in range(period):
    idx = np.arange(p + period, S - 1, period)      # skip the first layer (ring warm-up)
    print("slot %d: wait %8.0f (min %6.0f max %6.0f)  work %8.0f (min %6.0f max %6.0f)  publish %6.0f  stage-to-stage %8.0f" % (
        p, wait[:, idx].mean(), wait[:, idx].mean(1).min(), wait[:, idx].mean(1).max(),
        work[:, idx].mean(), work[:, idx].mean(1).min(), work[:, idx].mean(1).max(), pub[:, idx].mean(), tot[:, idx].mean()))
print("whole launch, CTA 0: %.0f cycles" % (t[0, -1, 3] - t[0, 0, 0]))
