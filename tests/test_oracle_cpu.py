"""CPU suite: pins the oracle (oracle/irfinder_cpu) on a hand-derived micro fixture and on frozen golden outputs,
and checks order-independence.  The mounted reference ships no golden vectors, so these pins
are first-principles derivations + a frozen regression, and parity with real IRFinder stays "unpinned"."""
import os
import random
import subprocess

import pytest

import micro_fixture
import micro_fixture2
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


def _check_rule_probe(ir_text, which):
    """The hand-derived cells of one alternative rule, and that every other cell kept its default value."""
    base = micro_fixture2.parse_ir(micro_fixture2.EXPECT["IRFinder-IR-nondir.txt"])
    got = micro_fixture2.parse_ir(ir_text)
    cells = micro_fixture2.RULE_PROBES[which][1]
    for intron, row in base.items():
        for col, dflt in row.items():
            want = cells.get(intron, {}).get(col, dflt)
            assert got[intron][col] == want, (which, intron, col, got[intron][col], want)


def test_rule_probe_fixture_by_hand(built, tmp_path):
    """tests/micro_fixture2.py: seven introns whose cells are decided by the Appendix-C rules (flank 4 vs 5, span point exactly at
    start+5 / end-4, percentile ties, edge windows, one intron per warning label), every cell derived by hand; then each rule is set
    to its alternative (IRF_RULE_*) and exactly the hand-derived cells change."""
    p = micro_fixture2.write(str(tmp_path / "m"))
    for key in ("collated", "sorted"):        # all reads are single-end, so the record order does not matter to the oracle either
        out = run_oracle(built, p, str(tmp_path / ("o_" + key)), bam=p[key])
        for name, want in micro_fixture2.EXPECT.items():
            assert open(os.path.join(str(tmp_path / ("o_" + key)), name)).read() == want, (key, name)
        st = _stats(out)
        for k, v in micro_fixture2.EXPECT_STATS.items():
            assert st[k] == v, (key, k)
    for which, (rules, _) in micro_fixture2.RULE_PROBES.items():
        d = str(tmp_path / ("r_" + which))
        run_oracle(built, p, d, env={"IRF_RULE_" + k.upper(): str(v) for k, v in rules.items()})
        _check_rule_probe(open(os.path.join(d, "IRFinder-IR-nondir.txt")).read(), which)


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


def _int_columns(path, key_cols, val_cols):
    rows = {}
    with open(path) as f:
        next(f)
        for line in f:
            c = line.rstrip("\n").split("\t")
            rows[tuple(c[i] for i in key_cols)] = [int(c[i]) for i in val_cols]
    return rows


def test_integer_tables_are_additive_over_fragments(built, tmp_path):
    """Linearity: every integer table is a sum over fragments, so splitting the templates of a BAM into two files must give
    tables that add up to the whole file's (junction, span-point, chromosome and ROI counts, and the read counters)."""
    from irfinder_b200 import synth
    meta = synth.generate(str(tmp_path / "w"), genome="tiny", introns=150, pairs=5000, readlen=100, seed=24, chunks=2)
    refs, recs = read_records(inflate_file(meta["collated"]))
    groups, i = [], 0
    while i < len(recs):
        if i + 1 < len(recs) and recs[i][3] == recs[i + 1][3]:
            groups.append(recs[i:i + 2]); i += 2
        else:
            groups.append(recs[i:i + 1]); i += 1
    rnd = random.Random(9)
    parts = ([], [])
    for g in groups:
        parts[rnd.randrange(2)].extend(r[6] for r in g)
    outs, stats = [], []
    for k, rs in enumerate(parts):
        bam = str(tmp_path / f"part{k}.bam")
        write_bam(bam, refs, rs, level=1)
        stats.append(_stats(run_oracle(built, meta, str(tmp_path / f"o{k}"), bam=bam)))
        outs.append(str(tmp_path / f"o{k}"))
    whole = _stats(run_oracle(built, meta, str(tmp_path / "all")))
    for key in ("Total reads processed", "Total nucleotides", "Total singles processed", "Total pairs processed", "Short pairs", "Intersect pairs", "Long pairs", "Skipped reads", "Error reads"):
        assert int(whole[key]) == int(stats[0][key]) + int(stats[1][key]), key
    for name, keys, vals in (("IRFinder-JuncCount.txt", (0, 1, 2), (4, 5, 6)), ("IRFinder-SpansPoint.txt", (0, 1), (2, 3, 4)),
                             ("IRFinder-ChrCoverage.txt", (0,), (1, 2, 3)), ("IRFinder-ROI.txt", (0,), (1, 2, 3))):
        a, b, w = (_int_columns(os.path.join(d, name), keys, vals) for d in (outs[0], outs[1], str(tmp_path / "all")))
        assert set(w) == set(a) | set(b), name          # a novel junction appears in the whole iff it appears in a part
        for k, v in w.items():
            va, vb = a.get(k, [0] * len(v)), b.get(k, [0] * len(v))
            assert v == [x + y for x, y in zip(va, vb)], (name, k)
    assert int(whole["Total pairs processed"]) > 1000
