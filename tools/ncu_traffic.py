#!/usr/bin/env python
"""DRAM traffic of the decode GEMV launches of ONE step, from an ncu --set full capture (tools/gpu_ncu_decode.sh):
   ncu_traffic.py <rep> <n_mulmat nodes per step> [out.json] [layers]
(with `layers`: the capture holds ONE layer's GEMV launches followed by the output matrix's; the step total is layers x the layer + the output)
writes {"dram_bytes_per_mulmat": (sum over the captured GEMV launches of dram__bytes_read.sum + dram__bytes_write.sum) / n_mulmat, ...}
for bench.py's roofline.traffic, and prints a per-kernel table (launches, mean us, DRAM MB per launch, DRAM GB/s inside the kernel)."""
import collections, csv, io, json, subprocess, sys

rep, n_mulmat = sys.argv[1], int(sys.argv[2])
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {k: hdr.index(k) for k in ("Kernel Name", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum")}


def val(r, k):
    v = float(r[col[k]].replace(",", ""))
    u = units[col[k]]
    return v * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}.get(u, 1.0)


agg = collections.OrderedDict()
tot = 0.0
layers = int(sys.argv[4]) if len(sys.argv) > 4 else 0
body = rows[2:]
for n_r, r in enumerate(body):
    k = (r[col["Kernel Name"]].split("(")[0][:60], r[col["Grid Size"]])
    b = val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
    a = agg.setdefault(k, [0, 0.0, 0.0])
    a[0] += 1; a[1] += val(r, "gpu__time_duration.sum"); a[2] += b
    tot += b * (layers if layers and n_r + 1 < len(body) else 1)
print("%-62s %-14s %5s %9s %11s %9s" % ("kernel", "grid", "n", "mean_us", "MB/launch", "GB/s"))
for (k, g), (n, us, b) in agg.items():
    print("%-62s %-14s %5d %9.2f %11.3f %9.1f" % (k, g, n, us / n, b / n / 1e6, b / us / 1e3))
out = {"dram_bytes_per_mulmat": int(tot / n_mulmat), "dram_bytes_step": int(tot), "gemv_launches_captured": sum(v[0] for v in agg.values()), "scaled_by_layers": layers,
       "n_mulmat_nodes": n_mulmat, "source": rep.split("/")[-1],
       "commit": subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()}
print(json.dumps(out))
if len(sys.argv) > 3:
    json.dump(out, open(sys.argv[3], "w"), indent=1)
