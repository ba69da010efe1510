"""CPU suite: pins the oracle (oracle/irfinder_cpu) on a hand-derived micro fixture and on frozen golden outputs,
and checks order-independence.  The mounted reference ships no golden vectors, so these pins
are first-principles derivations + a frozen regression, and parity with real IRFinder stays "unpinned"."""
import os
import random
import subprocess

import pytest

import micro_fixture
from bamutil import inflate_file, read_records, write_bam
from conftest import OUT_FILES, ROOT, run_oracle

GOLDEN = os.path.join(ROOT, "tests", "golden")


def _stats(text):
    return {l.split("\t")[0]: l.split("\t")[1] for l in text.strip().split("\n") if "\t" in l}


def test_micro_fixture_by_hand(built, tmp_path):
    p = micro_fixture.write(str(tmp_path / "m"))
    out = run_oracle(built, p, str(tmp_path / "o"))
    for name, want in micro_fixture.EXPECT.items():
        got = open(os.path.join(str(tmp_path / "o"), name)).read()
        assert got == want, name
    assert not os.path.exists(str(tmp_path / "o" / "IRFinder-IR-dir.txt"))
    st = _stats(out)
    for k, v in micro_fixture.EXPECT_STATS.items():
        assert st[k] == v, k
    assert st["Directionality:"] == "0"


def test_inflated_stdin_equals_bgzf_file(built, tmp_path):
    p = micro_fixture.write(str(tmp_path / "m"))
    raw = inflate_file(p["collated"])
    rp = str(tmp_path / "raw.bam")
    open(rp, "wb").write(raw)
    run_oracle(built, p, str(tmp_path / "a"))
    run_oracle(built, p, str(tmp_path / "b"), bam=rp)
    for name in micro_fixture.EXPECT:
        assert open(str(tmp_path / "a" / name)).read() == open(str(tmp_path / "b" / name)).read()


def test_golden_regression(built, tmp_path):
    """Seeded tiny workload -> outputs frozen under tests/golden/ by tests/golden/make_golden.py."""
    from irfinder_b200 import synth
    meta = synth.generate(str(tmp_path / "w"), genome="tiny", introns=120, pairs=4000, readlen=100, seed=21, chunks=3)
    run_oracle(built, meta, str(tmp_path / "o"))
    for name in OUT_FILES:
        g = os.path.join(GOLDEN, "tiny_seed21", name)
        o = str(tmp_path / "o" / name)
        assert os.path.exists(g) == os.path.exists(o), name
        if os.path.exists(g):
            assert open(g, "rb").read() == open(o, "rb").read(), name


def test_counts_do_not_depend_on_pair_order(built, tmp_path):
    from irfinder_b200 import synth
    meta = synth.generate(str(tmp_path / "w"), genome="tiny", introns=100, pairs=2500, readlen=100, seed=22, chunks=2)
    refs, recs = read_records(inflate_file(meta["collated"]))
    groups, i = [], 0
    while i < len(recs):   # keep mates adjacent, shuffle the templates
        if i + 1 < len(recs) and recs[i][3] == recs[i + 1][3]:
            groups.append(recs[i:i + 2]); i += 2
        else:
            groups.append(recs[i:i + 1]); i += 1
    random.Random(5).shuffle(groups)
    shuffled = str(tmp_path / "shuffled.bam")
    write_bam(shuffled, refs, [r[6] for g in groups for r in g], level=1)
    run_oracle(built, meta, str(tmp_path / "a"))
    run_oracle(built, meta, str(tmp_path / "b"), bam=shuffled)
    for name in OUT_FILES:
        pa = str(tmp_path / "a" / name)
        if os.path.exists(pa):
            assert open(pa, "rb").read() == open(str(tmp_path / "b" / name), "rb").read(), name


def test_generator_sorted_and_collated_hold_the_same_records(built, tmp_path):
    from irfinder_b200 import synth
    meta = synth.generate(str(tmp_path / "w"), genome="tiny", introns=80, pairs=1500, readlen=100, seed=23, chunks=4)
    _, a = read_records(inflate_file(meta["sorted"]))
    _, b = read_records(inflate_file(meta["collated"]))
    assert sorted(r[6] for r in a) == sorted(r[6] for r in b)
    keys = [((r[0] & 0xFFFFFFFF), r[1]) for r in a]
    assert keys == sorted(keys), "sorted.bam must be coordinate-sorted"
    assert open(meta["sorted"], "rb").read()[-28:] == open(meta["collated"], "rb").read()[-28:]   # BGZF EOF marker
    assert meta["records"] == len(a)


def test_oracle_rejects_truncated_stream(built, tmp_path):
    p = micro_fixture.write(str(tmp_path / "m"))
    raw = inflate_file(p["collated"])
    rp = str(tmp_path / "cut.bam")
    open(rp, "wb").write(raw[: len(raw) - 37])
    os.makedirs(str(tmp_path / "o"))
    r = subprocess.run([built["oracle"], str(tmp_path / "o"), p["cover"], p["sj"], p["rc"], p["roi"], rp], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert r.returncode != 0
