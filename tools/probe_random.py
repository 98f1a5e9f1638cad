#!/usr/bin/env python3
"""Random-access ceiling of the device: k_random_gather over table sizes and chunk sizes (GB/s)."""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "desamba-so_b200"))
import pydsb

cuda = pydsb.load_cuda_layer()
for gib in (0.25, 1, 2, 8, 32):
    row = []
    for chunk in (32, 64, 128):
        t = C.c_float(0)
        rc = cuda.dsb_random_access_probe(0, C.c_uint64(int(gib * (1 << 30))), C.c_uint32(chunk), C.c_uint64(1 << 26), 3, C.byref(t))
        if rc != 0:
            raise SystemExit(cuda.dsb_last_error().decode())
        row.append("%dB %.0f GB/s (%.1f G acc/s)" % (chunk, (1 << 26) * chunk / (t.value * 1e-3) / 1e9, (1 << 26) / (t.value * 1e-3) / 1e9))
    print("table %5.2f GiB: " % gib + "  ".join(row), flush=True)
