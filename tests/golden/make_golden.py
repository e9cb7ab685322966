#!/usr/bin/env python
"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref/scalar, built by
oracle/Makefile.ref from /root/reference) on seeded inputs.  Run in the build container:
    make -f oracle/Makefile.ref VARIANT=scalar && python tests/golden/make_golden.py
The fixtures pin oracle/ggml_oracle.c (tests/test_oracle.py) and the CUDA path (tests/test_gpu_parity.py)
on machines where /root/reference and oracle/_ref do not exist."""
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import pyref  # noqa: E402

Q4_0, Q8_0, Q8_1, Q4_K, Q6_K, Q8_K = 2, 8, 9, 12, 14, 15
R = pyref.Ref("scalar")
rng = np.random.default_rng(20251104)
out = {}

# activation quantisers: quantize_row_q8_{0,1,K}_ref on rows with edge cases
x = rng.uniform(-1, 1, (6, 512)).astype(np.float32)
x[1, :32] = 0.0                       # all-zero block: d = 0, id = 0
x[2] *= 1e-6                          # tiny values (fp16 subnormal scale)
x[3] *= 300.0                         # large values
x[4, ::2] = 0.5                       # ties in rounding
x[5] = rng.standard_normal(512).astype(np.float32)
out["act_x"] = x
for t, n in ((Q8_0, "q8_0"), (Q8_1, "q8_1"), (Q8_K, "q8_K")):
    out["act_" + n] = R.quantize_row_ref(t, x)

# weights via ggml_quantize_chunk (as test-backend-ops does), dequantize, vec_dot, full MUL_MAT on ggml-cpu
M, K, N = 24, 512, 3
wf = rng.uniform(-1, 1, (M, K)).astype(np.float32)
xs = rng.uniform(-1, 1, (N, K)).astype(np.float32)
out["mm_wf"], out["mm_x"] = wf, xs
for t, n in ((Q4_0, "q4_0"), (Q8_0, "q8_0"), (Q4_K, "q4_K"), (Q6_K, "q6_K")):
    w = R.quantize_chunk(t, wf)
    out["w_" + n] = w
    out["deq_" + n] = R.dequantize(t, w, K)
    out["mm_" + n] = R.mul_mat(t, w, xs)          # [N, M], scalar ggml-cpu: bit-exact target of the oracle
a = rng.standard_normal((2, 3, 4, 64)).astype(np.float32)
b = rng.standard_normal((1, 3, 1, 64)).astype(np.float32)
out["add_a"], out["add_b"], out["add_out"] = a, b, R.add(a, b)
np.savez_compressed(os.path.join(HERE, "ref_scalar.npz"), **out)
print("wrote", os.path.join(HERE, "ref_scalar.npz"), {k: v.shape for k, v in out.items()})

# ---- the ops ggml-hexagon lists but leaves as stubs: RMS_NORM, SOFT_MAX, POOL_2D on ggml-cpu
ops = {}
rng = np.random.default_rng(20251105)
x = rng.standard_normal((3, 4, 5, 64)).astype(np.float32)
x[0, 0, 0] = 0.0                                    # an all-zero row: scale = 1 / sqrt(eps)
x[0, 0, 1] *= 1e4
ops["rms_x"] = x
ops["rms_out_eps1e-6"] = pyref.ref_rms_norm(R, x, 1e-6)
ops["rms_out_eps0"] = pyref.ref_rms_norm(R, x[1:], 0.0)
xl = rng.standard_normal((2, 4096)).astype(np.float32)
ops["rms_xl"], ops["rms_outl"] = xl, pyref.ref_rms_norm(R, xl, 1e-5)
s = rng.standard_normal((2, 8, 5, 77)).astype(np.float32) * 3
m = rng.standard_normal((5, 77)).astype(np.float32)
m[m > 1.0] = -np.inf                                # causal-mask style -inf entries
ops["sm_x"], ops["sm_mask"] = s, m
ops["sm_plain"] = pyref.ref_soft_max(R, s, None, 1.0, 0.0)
ops["sm_mask_scale"] = pyref.ref_soft_max(R, s, m, 0.1, 0.0)
ops["sm_alibi"] = pyref.ref_soft_max(R, s, m, 0.1, 8.0)
sl = rng.standard_normal((1, 1, 3, 4100)).astype(np.float32)
ops["sm_xl"], ops["sm_outl"] = sl, pyref.ref_soft_max(R, sl, None, 0.125, 0.0)
p = rng.standard_normal((2, 3, 10, 11)).astype(np.float32)
ops["pool_x"] = p
for name, (op, k0, k1, s0, s1, p0, p1) in {"max_3311": (0, 3, 3, 1, 1, 1, 1), "avg_3322": (1, 3, 3, 2, 2, 1, 1), "avg_2321": (1, 2, 3, 2, 1, 0, 1),
                                          "max_1122": (0, 1, 1, 2, 2, 0, 0)}.items():
    ops["pool_" + name] = pyref.ref_pool_2d(R, p, op, k0, k1, s0, s1, p0, p1)
    ops["poolargs_" + name] = np.array([op, k0, k1, s0, s1, p0, p1], dtype=np.int32)
np.savez_compressed(os.path.join(HERE, "ref_scalar_ops.npz"), **ops)
print("wrote", os.path.join(HERE, "ref_scalar_ops.npz"), {k: v.shape for k, v in ops.items()})
