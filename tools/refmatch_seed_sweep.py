#!/usr/bin/env python3
"""Runs the reference pins of tests/test_oracle_refmatch.py (restatement vs the reference's compiled ORBmatcher.cc / Frame.cc / MapPoint.cc /
KannalaBrandt8.cpp / DBoW2) on OTHER synthetic worlds: every `seed=N` of the test file is shifted by an offset in a temporary copy.
CPU only; needs oracle/_ref (i.e. /root/reference at build time).

    python tools/refmatch_seed_sweep.py [offset ...]        default offsets: 1000 2000 3000 4000

A failure of a sanity bound (`nm > 10` on a world that happens to give few matches) is not a parity failure: read the assertion."""
import os
import re
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
offsets = [int(a) for a in sys.argv[1:]] or [1000, 2000, 3000, 4000]
tmp = tempfile.mkdtemp(prefix="refmatch_seeds_")
for f in os.listdir(os.path.join(ROOT, "tests")):
    if f.endswith(".py"):
        shutil.copy(os.path.join(ROOT, "tests", f), tmp)
src = open(os.path.join(ROOT, "tests", "test_oracle_refmatch.py")).read()
files = []
for off in offsets:
    path = os.path.join(tmp, "test_shifted_%d.py" % off)
    open(path, "w").write(re.sub(r"seed=(\d+)", lambda m: "seed=%d" % (int(m.group(1)) + off), src))
    files.append(path)
env = dict(os.environ, PYTHONPATH=ROOT)
rc = subprocess.call([sys.executable, "-m", "pytest", "-q", "-p", "no:cacheprovider", "--rootdir=" + tmp] + files, env=env, cwd=ROOT)
shutil.rmtree(tmp, ignore_errors=True)
sys.exit(rc)
