"""In-tree builds: the CUDA library (sm_100a), the `irfinder` CLI shim, the workload generator and the test oracle."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(cmd, cwd):
    r = subprocess.run(cmd, cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"build failed: {' '.join(cmd)}\n{r.stdout[-4000:]}")
    return r.stdout


def build_library():
    _run(["make", "-j4"], os.path.join(ROOT, "irfinder_b200", "csrc"))
    return os.path.join(ROOT, "irfinder_b200", "libirfinder_b200.so")


def build_tools():
    os.makedirs(os.path.join(ROOT, "bin"), exist_ok=True)
    out = os.path.join(ROOT, "tools", "irf_synth")
    src = os.path.join(ROOT, "tools", "irf_synth.cpp")
    if not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
        _run(["g++", "-O3", "-std=c++17", "-pthread", "-o", out, src, "-lz"], ROOT)
    return out


def build_oracle():
    """Test infrastructure only (see oracle/irfinder_cpu.cpp header): building the checker is not using it."""
    _run(["make"], os.path.join(ROOT, "oracle"))
    return os.path.join(ROOT, "oracle", "irfinder_cpu")
