import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

OUT_FILES = ["IRFinder-IR-nondir.txt", "IRFinder-IR-dir.txt", "IRFinder-JuncCount.txt", "IRFinder-SpansPoint.txt", "IRFinder-ChrCoverage.txt", "IRFinder-ROI.txt"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `pytest -m gpu`)")


@pytest.fixture(scope="session")
def built():
    """Builds (or re-uses) the generator, the oracle and — when nvcc is present — the CUDA library."""
    from irfinder_b200 import build
    paths = dict(synth=build.build_tools(), oracle=build.build_oracle())
    lib = os.path.join(ROOT, "irfinder_b200", "libirfinder_b200.so")
    if not os.path.exists(lib):
        build.build_library()
    paths["lib"] = lib
    return paths


def run_oracle(built, meta, outdir, bam=None, dump=None, env=None, roi=True, depth=None, depth_digest=None):
    """Runs the CPU restatement (tests are the only place allowed to)."""
    os.makedirs(outdir, exist_ok=True)
    cmd = [built["oracle"], outdir, meta["cover"], meta["sj"], meta["rc"], meta["roi"] if roi else "NULL", bam or meta["collated"]]
    if dump:
        cmd += ["--dump-fragments", dump]
    if depth:
        cmd += ["--dump-depth", depth]
    if depth_digest:
        cmd += ["--depth-digest", depth_digest]
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=e)
    assert r.returncode == 0, r.stderr
    return r.stdout


def assert_same_outputs(dir_a, dir_b):
    for name in OUT_FILES:
        pa, pb = os.path.join(dir_a, name), os.path.join(dir_b, name)
        assert os.path.exists(pa) == os.path.exists(pb), f"{name}: present in one output directory only"
        if os.path.exists(pa):
            with open(pa, "rb") as fa, open(pb, "rb") as fb:
                a, b = fa.read(), fb.read()
            if a != b:
                la, lb = a.split(b"\n"), b.split(b"\n")
                for i, (x, y) in enumerate(zip(la, lb)):
                    if x != y:
                        raise AssertionError(f"{name} differs at line {i + 1}:\n  oracle: {x[:300]!r}\n  gpu:    {y[:300]!r}")
                raise AssertionError(f"{name} differs in length: {len(la)} vs {len(lb)} lines")


def depth_digest(v):
    """The oracle's --depth-digest of an int32 sequence: (n, sum, sum v[i]*(i+1), sum v[i]*(i+1)^2, max), sums mod 2^64."""
    import numpy as np
    n, s, w1, w2, mx = len(v), 0, 0, 0, 0
    M = (1 << 64) - 1
    for lo in range(0, n, 1 << 24):
        x = np.asarray(v[lo:lo + (1 << 24)]).astype(np.int64).view(np.uint64)
        i = np.arange(lo + 1, lo + 1 + len(x), dtype=np.uint64)
        with np.errstate(over="ignore"):
            s = (s + int(x.sum(dtype=np.uint64))) & M
            w1 = (w1 + int((x * i).sum(dtype=np.uint64))) & M
            w2 = (w2 + int((x * (i * i)).sum(dtype=np.uint64))) & M
        mx = max(mx, int(np.asarray(v[lo:lo + (1 << 24)]).max(initial=0)))
    return n, s, w1, w2, mx
