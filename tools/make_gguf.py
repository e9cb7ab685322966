#!/usr/bin/env python
"""Write a synthetic Llama-architecture GGUF (random-init quantised weights, dummy SPM vocabulary) so that the reference's own
llama-bench can run tg128 / pp512 through the ggml scheduler on libggml-b200.so.  There are no model files on
the box and no network: the shapes are the public Llama-2-7B ones, the values are random blocks with scales sized so that the
activations stay finite through 32 layers.

The container format is restated from its definition (ggml/src/gguf.cpp:218-224 type table, :340-520 reader, docs/gguf.md is
not in the tree): magic "GGUF", u32 version 3, i64 n_tensors, i64 n_kv; each kv = string key, i32 type, value; each tensor info =
string name, u32 n_dims, i64 ne[], i32 ggml_type, u64 offset; then the data section aligned to general.alignment (32).
Key names: src/llama-arch.cpp:76-180 (LLM_KV_*), tensor names :200-230 (LLM_TENSOR_NAMES for LLM_ARCH_LLAMA).

usage: make_gguf.py OUT.gguf [--preset 7b|8b|tiny] [--type q4_0|q8_0|q4_K|q6_K] [--kmix] [--layers N]
(8b = the Llama-3-8B shapes of BASELINE.json configs[3]: 8 KV heads, 14336-wide FFN, 128256-entry vocabulary; --kmix = attn_v and ffn_down in
q6_K, the rest in --type, as a Q4_K_M file is laid out)
"""
import argparse
import struct
import sys

import numpy as np

GGUF_U32, GGUF_I32, GGUF_F32, GGUF_STR, GGUF_ARR = 4, 5, 6, 8, 9
GGML_F32, GGML_Q4_0, GGML_Q8_0, GGML_Q4_K, GGML_Q6_K = 0, 2, 8, 12, 14
BLOCK = {GGML_Q4_0: (32, 18), GGML_Q8_0: (32, 34), GGML_Q4_K: (256, 144), GGML_Q6_K: (256, 210)}   # elements, bytes (ggml-common.h:167-326)
TYPES = {"q4_0": GGML_Q4_0, "q8_0": GGML_Q8_0, "q4_K": GGML_Q4_K, "q6_K": GGML_Q6_K}
FTYPE = {GGML_Q4_0: 2, GGML_Q8_0: 7, GGML_Q4_K: 15, GGML_Q6_K: 18}      # llama_ftype (include/llama.h LLAMA_FTYPE_MOSTLY_*)
ALIGN = 32


def s(x):
    b = x.encode()
    return struct.pack("<Q", len(b)) + b


def kv(key, typ, val):
    out = s(key) + struct.pack("<i", typ)
    if typ == GGUF_U32:
        return out + struct.pack("<I", val)
    if typ == GGUF_I32:
        return out + struct.pack("<i", val)
    if typ == GGUF_F32:
        return out + struct.pack("<f", val)
    if typ == GGUF_STR:
        return out + s(val)
    raise ValueError(typ)


def kv_arr(key, etyp, vals):
    out = s(key) + struct.pack("<i", GGUF_ARR) + struct.pack("<iQ", etyp, len(vals))
    if etyp == GGUF_STR:
        return out + b"".join(s(v) for v in vals)
    if etyp == GGUF_F32:
        return out + np.asarray(vals, np.float32).tobytes()
    if etyp == GGUF_I32:
        return out + np.asarray(vals, np.int32).tobytes()
    raise ValueError(etyp)


def random_blocks(t, rows, K, rng, sigma):
    """rows x K weights as native ggml blocks: random quants, fp16 super-scales chosen so that one weight has std ~ sigma"""
    elems, nbytes = BLOCK[t]
    nb = rows * (K // elems)
    blk = rng.integers(0, 256, (nb, nbytes), dtype=np.uint8)
    h = lambda v: np.full(nb, v, np.float16).view(np.uint8).reshape(nb, 2)
    if t == GGML_Q4_0:                                   # {half d; u8 qs[16]}: w = d*(q-8), q uniform 0..15 -> std 4.6 d
        blk[:, 0:2] = h(sigma / 4.6)
    elif t == GGML_Q8_0:                                 # {half d; i8 qs[32]}: std 74 d
        blk[:, 0:2] = h(sigma / 74.0)
    elif t == GGML_Q4_K:                                 # {half d, dmin; u8 scales[12]; u8 qs[128]}: w = d*sc*q - dmin*m, sc,m 0..63
        blk[:, 0:2] = h(sigma / (4.6 * 36.0))
        blk[:, 2:4] = h(sigma / (4.6 * 36.0) * 7.5)      # centres q (mean 7.5) when m ~ sc
    elif t == GGML_Q6_K:                                 # {u8 ql[128], qh[64]; i8 scales[16]; half d}: w = d*sc*(q-32), sc -128..127
        blk[:, 208:210] = h(sigma / (18.5 * 74.0))
    return blk


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--preset", default="7b", choices=["7b", "8b", "tiny"])
    ap.add_argument("--kmix", action="store_true", help="attn_v and ffn_down in q6_K (Q4_K_M-style mix)")
    ap.add_argument("--type", default="q4_0", choices=sorted(TYPES))
    ap.add_argument("--out-type", default="q6_K", choices=sorted(TYPES), help="type of output.weight (llama-quantize keeps it at q6_K in a Q4_0 file)")
    ap.add_argument("--rope-freqs", action="store_true", help="add rope_freqs.weight (Llama-3.1-style frequency factors: ROPE gets a third operand)")
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--seed", type=int, default=1234)
    a = ap.parse_args()
    freq_base = 10000.0
    if a.preset == "7b":
        n_embd, n_ff, n_head, n_head_kv, n_layer, n_vocab, n_ctx = 4096, 11008, 32, 32, 32, 32000, 4096
    elif a.preset == "8b":
        n_embd, n_ff, n_head, n_head_kv, n_layer, n_vocab, n_ctx, freq_base = 4096, 14336, 32, 8, 32, 128256, 8192, 500000.0
    else:
        n_embd, n_ff, n_head, n_head_kv, n_layer, n_vocab, n_ctx = 256, 512, 4, 4, 2, 512, 512
    if a.layers:
        n_layer = a.layers
    t, t_out = TYPES[a.type], TYPES[a.out_type]
    if a.preset == "tiny":
        t_out = t if BLOCK[t][0] == 32 else t_out
    rng = np.random.default_rng(a.seed)

    # vocabulary in the shape of the Llama SPM one: <unk>, <s>, </s>, 256 byte tokens, then plain pieces (llama-vocab.cpp reads
    # tokens / scores / token_type and looks the byte tokens up by their "<0xNN>" text)
    tokens = ["<unk>", "<s>", "</s>"] + ["<0x%02X>" % i for i in range(256)] + ["▁t%d" % i for i in range(n_vocab - 259)]
    ttype = [2, 3, 3] + [6] * 256 + [1] * (n_vocab - 259)              # UNKNOWN, CONTROL, CONTROL, BYTE, NORMAL
    scores = [0.0] * 259 + [-float(i) for i in range(n_vocab - 259)]

    kvs = [
        kv("general.architecture", GGUF_STR, "llama"),
        kv("general.name", GGUF_STR, "synthetic-llama-%s-%s" % (a.preset, a.type)),
        kv("general.file_type", GGUF_U32, FTYPE[t]),
        kv("general.alignment", GGUF_U32, ALIGN),
        kv("llama.vocab_size", GGUF_U32, n_vocab),
        kv("llama.context_length", GGUF_U32, n_ctx),
        kv("llama.embedding_length", GGUF_U32, n_embd),
        kv("llama.block_count", GGUF_U32, n_layer),
        kv("llama.feed_forward_length", GGUF_U32, n_ff),
        kv("llama.rope.dimension_count", GGUF_U32, n_embd // n_head),
        kv("llama.attention.head_count", GGUF_U32, n_head),
        kv("llama.attention.head_count_kv", GGUF_U32, n_head_kv),
        kv("llama.attention.layer_norm_rms_epsilon", GGUF_F32, 1e-5),
        kv("llama.rope.freq_base", GGUF_F32, freq_base),
        kv("tokenizer.ggml.model", GGUF_STR, "llama"),
        kv_arr("tokenizer.ggml.tokens", GGUF_STR, tokens),
        kv_arr("tokenizer.ggml.scores", GGUF_F32, scores),
        kv_arr("tokenizer.ggml.token_type", GGUF_I32, ttype),
        kv("tokenizer.ggml.bos_token_id", GGUF_U32, 1),
        kv("tokenizer.ggml.eos_token_id", GGUF_U32, 2),
        kv("tokenizer.ggml.unknown_token_id", GGUF_U32, 0),
    ]

    # tensor list: (name, ne (ggml order: ne[0] = K contiguous), type, generator)
    sig = lambda K: 1.0 / np.sqrt(K)
    tensors = [("token_embd.weight", (n_embd, n_vocab), t, lambda: random_blocks(t, n_vocab, n_embd, rng, 1.0))]
    n_kv_dim = n_embd // n_head * n_head_kv
    # one layer's random blocks are drawn once and written for every layer: the bytes are distinct tensors in the file and in HBM,
    # identical content does not change what the kernels do, and it keeps the 3.8 GB file to seconds of generation
    cache = {}

    t6 = GGML_Q6_K if a.kmix else t

    def lw(key, rows, K, ty=None):
        ty = t if ty is None else ty

        def gen():
            if key not in cache:
                cache[key] = random_blocks(ty, rows, K, rng, sig(K))
            return cache[key]
        return gen
    ones = lambda: np.ones(n_embd, np.float32)
    for il in range(n_layer):
        p = "blk.%d." % il
        tensors += [
            (p + "attn_norm.weight", (n_embd,), GGML_F32, ones),
            (p + "attn_q.weight", (n_embd, n_embd), t, lw("q", n_embd, n_embd)),
            (p + "attn_k.weight", (n_embd, n_kv_dim), t, lw("k", n_kv_dim, n_embd)),
            (p + "attn_v.weight", (n_embd, n_kv_dim), t6, lw("v", n_kv_dim, n_embd, t6)),
            (p + "attn_output.weight", (n_embd, n_embd), t, lw("o", n_embd, n_embd)),
            (p + "ffn_norm.weight", (n_embd,), GGML_F32, ones),
            (p + "ffn_gate.weight", (n_embd, n_ff), t, lw("gate", n_ff, n_embd)),
            (p + "ffn_down.weight", (n_ff, n_embd), t6, lw("down", n_embd, n_ff, t6)),
            (p + "ffn_up.weight", (n_embd, n_ff), t, lw("up", n_ff, n_embd)),
        ]
    if a.rope_freqs:
        hd2 = n_embd // n_head // 2
        tensors.append(("rope_freqs.weight", (hd2,), GGML_F32, lambda: np.linspace(1.0, 8.0, hd2).astype(np.float32)))
    tensors += [("output_norm.weight", (n_embd,), GGML_F32, ones),
                ("output.weight", (n_embd, n_vocab), t_out, lambda: random_blocks(t_out, n_vocab, n_embd, rng, sig(n_embd)))]

    def nbytes(ne, ty):
        n = int(np.prod(ne))
        return n * 4 if ty == GGML_F32 else n // BLOCK[ty][0] * BLOCK[ty][1]
    pad = lambda n: (n + ALIGN - 1) // ALIGN * ALIGN
    infos, off = [], 0
    for name, ne, ty, _ in tensors:
        infos.append(s(name) + struct.pack("<I", len(ne)) + struct.pack("<%dq" % len(ne), *ne) + struct.pack("<iQ", ty, off))
        off += pad(nbytes(ne, ty))
    head = b"GGUF" + struct.pack("<Iqq", 3, len(tensors), len(kvs)) + b"".join(kvs) + b"".join(infos)
    with open(a.out, "wb") as f:
        f.write(head)
        f.write(b"\0" * (pad(len(head)) - len(head)))
        for name, ne, ty, gen in tensors:
            data = gen()
            n = nbytes(ne, ty)
            assert data.nbytes == n, (name, data.nbytes, n)
            f.write(memoryview(data).cast("B"))
            f.write(b"\0" * (pad(n) - n))
    print("wrote %s: %d tensors, %.2f GB" % (a.out, len(tensors), (pad(len(head)) + off) / 1e9), file=sys.stderr)


if __name__ == "__main__":
    main()
