#!/usr/bin/env python3
"""Random 32-byte gathers (k_random_gather) from tables of 16 MiB .. 1 GiB: how large a randomly probed table may be and
still be served by the L2 (the exist tables are 256 MiB .. 1 GiB; probing them in hash-range passes is only worth it if a
pass's slice of the table stays resident)."""
import ctypes as C
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "desamba-so_b200"))
import pydsb  # noqa: E402
cuda = pydsb.load_cuda_layer()
n = 1 << 27
for mib in (16, 32, 48, 64, 80, 96, 128, 192, 256, 512, 1024):
    t = C.c_float(0)
    rc = cuda.dsb_random_access_probe(0, C.c_uint64(mib << 20), C.c_uint32(32), C.c_uint64(n), 3, C.byref(t))
    if rc != 0:
        raise RuntimeError(cuda.dsb_last_error().decode())
    print("table %5d MiB: %6.1f G gathers/s (%.3f ms for 2^27 gathers)" % (mib, n / (t.value * 1e-3) / 1e9, t.value), flush=True)
