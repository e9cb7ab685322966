"""Host-side bookkeeping of the row split: which weight rows a rank owns and how the output
slices line up in the full vector.  Mirrors what ggml-cuda's split buffers do per device (ggml-cuda.cu:727-753), for
one process per GPU.  Used by bench.py and exercised on CPU with gloo in tests/test_rowsplit_host.py."""


def row_range(M, rank, world, align=1):
    """rows [lo, hi) of an M-row matrix owned by `rank`; equal slices (the all-gather and the peer stores need them equal)"""
    if M % (world * align):
        raise ValueError("M=%d does not split into %d equal slices of a multiple of %d rows" % (M, world, align))
    rows = M // world
    return rank * rows, (rank + 1) * rows


def slice_rows(raw, rank, world):
    """this rank's rows of a [M, row_bytes] weight matrix"""
    lo, hi = row_range(raw.shape[0], rank, world)
    return raw[lo:hi]


def dst_offset_bytes(M, rank, world, column=0):
    """byte offset of this rank's slice inside column `column` of the full f32 dst [N][M]"""
    lo, _ = row_range(M, rank, world)
    return 4 * (column * M + lo)
