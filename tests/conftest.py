"""Shared fixtures.  `-m "not gpu"` runs everywhere; `-m gpu` needs a B200 and calls through the C-ABI."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "desamba-so_b200")
sys.path.insert(0, PKG)
sys.path.insert(0, os.path.join(ROOT, "tools"))
GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); runs the sm_100a kernels through the C-ABI")


def _make(args):
    r = subprocess.run(["make"] + args, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    assert r.returncode == 0, r.stdout.decode()[-3000:]


@pytest.fixture(scope="session")
def built():
    """Everything the tests load: product libraries, host simulator, oracle (checker)."""
    if not os.path.exists(os.path.join(PKG, "lib", "libdesamba.so")):
        _make(["-C", PKG, "all"])
    if not os.path.exists(os.path.join(ROOT, "tests", "hostsim", "libdesamba_hostsim.so")):
        _make(["-C", PKG, "hostsim"])
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "liboracle_port.so")):
        _make(["-C", os.path.join(ROOT, "oracle"), "port"])
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "deSAMBA")) and os.path.isdir("/root/reference/src"):
        _make(["-C", os.path.join(ROOT, "oracle"), "ref"])
    return True


@pytest.fixture(scope="session")
def have_ref_binary(built):
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "deSAMBA"))


@pytest.fixture(scope="session")
def tiny(built, have_ref_binary):
    if not have_ref_binary:
        pytest.skip("oracle/_ref/deSAMBA (needed to build a test index) is absent")
    import dataset
    return dataset.build("tiny")


@pytest.fixture(scope="session")
def mixed(built, have_ref_binary):
    if not have_ref_binary:
        pytest.skip("oracle/_ref/deSAMBA (needed to build a test index) is absent")
    import dataset
    return dataset.build("mixed")


def sam_records(text):
    """{read name: [classification fields of each record]} -- QNAME FLAG RNAME POS MAPQ CIGAR AS."""
    d = {}
    if isinstance(text, bytes):
        text = text.decode()
    for line in text.split("\n"):
        if not line:
            continue
        f = line.split("\t")
        d.setdefault(f[0], []).append(tuple(f[:6] + f[11:12]))
    return d
