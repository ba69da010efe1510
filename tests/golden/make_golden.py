"""Regenerates tests/golden/tiny_seed21/*: the oracle's outputs on the seeded tiny workload used by
tests/test_oracle_cpu.py::test_golden_regression.  Run from the repo root:  python tests/golden/make_golden.py
(The mounted reference has no source or fixtures — SURVEY.md §0 — so these vectors freeze the in-repo CPU
restatement, reviewed against the hand-derived micro fixture; they do not come from IRFinder itself.)"""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from irfinder_b200 import build, synth  # noqa: E402

out = os.path.join(ROOT, "tests", "golden", "tiny_seed21")
os.makedirs(out, exist_ok=True)
oracle = build.build_oracle()
with tempfile.TemporaryDirectory() as tmp:
    meta = synth.generate(os.path.join(tmp, "w"), genome="tiny", introns=120, pairs=4000, readlen=100, seed=21, chunks=3)
    with open(os.path.join(out, "stdout.txt"), "w") as f:
        subprocess.run([oracle, out, meta["cover"], meta["sj"], meta["rc"], meta["roi"], meta["collated"]], check=True, stdout=f)
print("wrote", sorted(os.listdir(out)))
