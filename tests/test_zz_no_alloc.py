"""dsb_exist_batch_static: the exist-k-mer probe entry point that allocates nothing (include/dsb_cuda.h; no counterpart in
the reference).  Collected last on purpose.

Its kernel works on a file-scope `static __device__` array instead of cudaMalloc'd buffers; the result must equal the
existing entry point's (dsb_exist_batch) byte for byte -- the tolerance of every test of this integer path -- on two calls
in a row (the array keeps the previous call's contents), and sizes beyond the stated maximum must be rejected."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

STATIC_MAX = 4194304            # DSB_EXIST_STATIC_MAX of include/dsb_cuda.h


def test_exist_probes_without_allocation_equal_the_allocating_entry_point(built, tiny):
    import torch
    import pydsb
    ix = pydsb.DeviceIndex(tiny["idx"])
    k = ix.ek_len()
    rng = np.random.default_rng(5)
    # real read bases (so that a good share of the k-mers exist) followed by random ones
    seq = b"".join(open(tiny["reads"], "rb").read().split(b"\n")[1::4])
    lut = np.full(256, 1, dtype=np.uint8)
    for c, v in zip(b"ACGTacgt", (0, 1, 2, 3, 0, 1, 2, 3)):
        lut[c] = v
    real = lut[np.frombuffer(seq, dtype=np.uint8)]
    first = np.concatenate([real[:600000], rng.integers(0, 4, size=100000, dtype=np.uint8)])
    second = np.concatenate([rng.integers(0, 4, size=5000, dtype=np.uint8), real[600000:900000]])   # shorter: stale bytes behind it
    second[100:140] = 2                                                                                # a low-complexity stretch
    free0 = None
    for rep in range(2):                        # two rounds: every call after the first sees the previous call's scratch
        for codes in (first, second, first[:k], first):
            want, _ms = ix.exist(codes)
            got = ix.exist_static(codes)
            assert got.shape == want.shape
            assert (got == want).all(), (rep, len(codes), int((got != want).sum()))
        if rep == 0:
            torch.cuda.synchronize()
            free0 = torch.cuda.mem_get_info(0)[0]
    assert ix.exist(first)[0].mean() > 0.05     # the comparison is not vacuous
    # the static entry point itself allocates nothing: device memory after round 2 == after round 1
    for _ in range(5):
        ix.exist_static(second)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(0)[0] == free0
    # beyond the maximum: a clear error, nothing computed
    big = np.zeros(STATIC_MAX + 1, dtype=np.uint8)
    with pytest.raises(RuntimeError, match="exceed the static scratch"):
        ix.exist_static(big)
    with pytest.raises(RuntimeError, match="shorter than k"):
        ix.exist_static(first[:k - 1])
    # exactly the maximum still works and agrees
    edge = rng.integers(0, 4, size=STATIC_MAX, dtype=np.uint8)
    assert (ix.exist_static(edge) == ix.exist(edge)[0]).all()
    ix.close()
