"""CPU suite: the text writers of the library (csrc/refs.cpp is plain host C++) against printf("%g") and a hand-written table."""
import os
import subprocess

from conftest import ROOT


def test_cells_print_like_printf_g(tmp_path):
    exe = str(tmp_path / "format_check")
    cs = os.path.join(ROOT, "irfinder_b200", "csrc")
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", "-I", cs, "-I", "/usr/local/cuda/include", "-o", exe,
                    os.path.join(ROOT, "tests", "emu", "format_check.cpp"), os.path.join(cs, "refs.cpp")], check=True)
    r = subprocess.run([exe, "1000000"], stdout=subprocess.PIPE, text=True)
    assert r.returncode == 0 and " bad 0" in r.stdout, r.stdout
