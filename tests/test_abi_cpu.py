"""CPU suite: the C-ABI library loads, exports every symbol include/irfinder_b200.h declares, and refuses to run without a GPU."""
import ctypes
import os
import re
import subprocess

import pytest

from conftest import ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "irfinder_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(irf_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound(built):
    from irfinder_b200 import api
    lib = ctypes.CDLL(built["lib"])
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/irfinder_b200.h but not exported"
    assert sorted(api.SYMBOLS) == declared, "irfinder_b200/api.py must bind exactly the declared entry points"
    nm = subprocess.run(["nm", "-D", "--defined-only", built["lib"]], stdout=subprocess.PIPE, text=True).stdout
    exported = sorted(set(re.findall(r"\bT (irf_[a-z_0-9]+)\b", nm)))
    assert exported == declared, "the library must not export undeclared irf_* symbols"


def test_no_cpu_fallback_without_a_gpu(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from irfinder_b200 import IRFinderCore, IRFError
    with pytest.raises(IRFError):
        IRFinderCore(0)


def test_product_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under irfinder_b200/ or bin/ may include, import or execute it
    (build.py only compiles it for the tests and bench.py's CPU legs)."""
    pat = re.compile(r"oracle")
    for base in ("irfinder_b200/csrc", "irfinder_b200/cli", "include"):
        for dp, _, files in os.walk(os.path.join(ROOT, base)):
            for f in files:
                if f.endswith((".o", ".so")):
                    continue
                text = open(os.path.join(dp, f), errors="replace").read()
                hits = [l for l in text.split("\n") if pat.search(l) and not l.lstrip().startswith(("//", "*", "/*", "#"))]
                assert not hits, f"{base}/{f} mentions the oracle outside comments: {hits[:2]}"
    for f in ("api.py", "distributed.py", "__init__.py", "synth.py"):
        text = open(os.path.join(ROOT, "irfinder_b200", f)).read()
        assert "irfinder_cpu" not in text and "build_oracle" not in text, f


def test_library_links_no_zlib(built):
    """Inflate is a device kernel; the shipped library must not pull in zlib."""
    out = subprocess.run(["ldd", built["lib"]], stdout=subprocess.PIPE, text=True).stdout
    assert "libz." not in out


def test_cli_reports_usage_and_missing_gpu(built):
    """The `irfinder` drop-in prints the reference's usage line for a wrong argument count and, on a box without a CUDA
    device, fails with a clear message instead of falling back to anything."""
    import subprocess
    exe = os.path.join(ROOT, "bin", "irfinder")
    if not os.path.exists(exe):
        pytest.skip("CLI not built")
    r = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 1 and "Usage: irfinder <outputDirectory>" in r.stderr
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([exe, "/tmp", "a", "b", "c", "NULL", "x.bam"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        assert r.returncode == 2 and "cannot initialise CUDA device" in r.stderr and r.stdout == ""


def test_cli_device_list_is_parsed_before_cuda(built):
    """IRF_DEVICES (the in-process multi-device context): a list that cannot be parsed is refused with exit code 1 before any
    device is touched; well-formed lists ("0,1", ranges "0-3", repeats "0,0") get as far as device initialisation, which on a box
    without a GPU is the 'cannot initialise' failure that names the list."""
    exe = os.path.join(ROOT, "bin", "irfinder")
    if not os.path.exists(exe):
        pytest.skip("CLI not built")
    args = [exe, "/tmp", "a", "b", "c", "NULL", "x.bam"]
    for bad in ("0;1", "a", "3-1", "0,,1", "-1", "0-"):
        r = subprocess.run(args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=dict(os.environ, IRF_DEVICES=bad))
        assert r.returncode == 1 and "cannot parse IRF_DEVICES" in r.stderr, (bad, r.returncode, r.stderr)
    import torch
    if not torch.cuda.is_available():
        for good in ("0,1", "0-3", "0,0", "2"):
            r = subprocess.run(args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=dict(os.environ, IRF_DEVICES=good))
            assert r.returncode == 2 and "cannot initialise CUDA device" in r.stderr, (good, r.returncode, r.stderr)
