"""TEST INFRASTRUCTURE: the pydsb driver classes bound to the host simulator's libraries (tests/hostsim/*.so, the kernel
source compiled for the host).  The package itself (desamba-so_b200/pydsb.py) can only load the CUDA build."""
import os

import pydsb

SIM = os.path.dirname(os.path.abspath(__file__))
API = os.path.join(SIM, "libdesamba_hostsim.so")
CUDA = os.path.join(SIM, "libdsb_hostsim.so")


class Desamba(pydsb.Desamba):
    API_LIB = API
    CUDA_LIB = CUDA


class DeviceIndex(pydsb.DeviceIndex):
    API_LIB = API
    CUDA_LIB = CUDA


def load_cuda_layer():
    return pydsb.load_cuda_layer(CUDA)


def synthetic_fm_index(n_sym, **kw):
    return pydsb.synthetic_fm_index(n_sym, _libs=(API, CUDA), **kw)


def build_index(ref_fasta, out_dir, kmer_srt=None):
    return pydsb.build_index(ref_fasta, out_dir, kmer_srt, _cuda_lib=CUDA)
