#!/usr/bin/env python3
"""Regenerate tests/golden/ from the UNMODIFIED reference (needs oracle/_ref built from /root/reference).

  tests/golden/<set>.sam            reference `classify -t 1 -f SAM` output for the named synthetic set
  tests/golden/<set>.dropped.txt    reads the reference binary itself crashes on (excluded on both sides)
  tests/golden/<set>.meta.json      read / base counts and md5 of the SAM
  tests/golden/kat_tiny/*.bin       function-level known answers written by oracle/kat_harness.c
                                    (which #includes the reference's cly.c in place)
Everything is a pure function of the seeds in tools/dataset.py.
"""
import json
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import dataset  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    os.makedirs(GOLD, exist_ok=True)
    for name in ("tiny", "mixed"):
        shutil.rmtree(os.path.join(dataset.data_root(), name), ignore_errors=True)
        committed = os.path.join(GOLD, name + ".dropped.txt")
        if os.path.exists(committed):
            os.remove(committed)
        m = dataset.build(name)
        shutil.copy(m["golden"], os.path.join(GOLD, name + ".sam"))
        shutil.copy(os.path.join(m["dir"], "dropped.txt"), committed)
        json.dump({k: m[k] for k in ("name", "n_reads", "n_bases", "dropped", "golden_md5")},
                  open(os.path.join(GOLD, name + ".meta.json"), "w"), indent=1)
    m = dataset.build("tiny")
    kat = os.path.join(GOLD, "kat_tiny")
    os.makedirs(kat, exist_ok=True)
    subprocess.check_call([os.path.join(ROOT, "oracle", "_ref", "kat_harness"), m["idx"], m["reads"], kat, "12"])
    print("golden vectors written to", GOLD)


if __name__ == "__main__":
    main()
