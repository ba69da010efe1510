"""GPU parity: the CUDA path through the C ABI vs the CPU restatement (oracle/) on the same seeded inputs.

Integer tables must be bit-exact and the text tables byte-identical (the floating columns are printed with
6 significant digits, so byte identity is stricter than BASELINE.json's 1e-6 relative tolerance).
"""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, assert_same_outputs, run_oracle

pytestmark = pytest.mark.gpu


def _core(**opt):
    from irfinder_b200 import IRFinderCore
    return IRFinderCore(0, **opt)


def _gen(tmp, name, **kw):
    from irfinder_b200 import synth
    return synth.generate(os.path.join(str(tmp), name), **kw)


def _gpu_run(meta, outdir, bam=None, roi=True, **opt):
    c = _core(**opt)
    c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"] if roi else None)
    c.run_bam(bam or meta["sorted"])
    c.write_outputs(outdir)
    return c


def _stats_from_stdout(text):
    return {l.split("\t")[0]: l.split("\t")[1] for l in text.strip().split("\n") if "\t" in l}


@pytest.fixture(scope="module")
def tiny(built, tmp_path_factory):
    tmp = tmp_path_factory.mktemp("tiny")
    meta = _gen(tmp, "w", genome="tiny", introns=300, pairs=30000, readlen=100, seed=11, chunks=7)
    meta["oracle_out"] = os.path.join(str(tmp), "oracle")
    meta["oracle_depth"] = os.path.join(str(tmp), "depth.i32")
    meta["oracle_depth_digest"] = os.path.join(str(tmp), "depth.digest")
    meta["oracle_stdout"] = run_oracle(built, meta, meta["oracle_out"], dump=os.path.join(str(tmp), "frags.txt"), depth=meta["oracle_depth"], depth_digest=meta["oracle_depth_digest"])
    meta["oracle_frags"] = os.path.join(str(tmp), "frags.txt")
    return meta


def test_sorted_bam_matches_oracle(tiny, tmp_path):
    c = _gpu_run(tiny, str(tmp_path / "gpu"))
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "gpu"))
    assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(tiny["oracle_stdout"])


def test_collated_bam_matches_oracle(tiny, tmp_path):
    c = _gpu_run(tiny, str(tmp_path / "gpu"), bam=tiny["collated"])
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "gpu"))
    assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(tiny["oracle_stdout"])


def test_fragment_blocks_stream(tiny, tmp_path):
    """FragmentBlocks values, element for element (order-free): chr, direction, readCount and every block."""
    c = _core(dump_fragments=1)
    c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
    c.run_bam(tiny["sorted"])
    w = c.table("fragments")
    got, i = [], 0
    while i < len(w):
        n, chr_, d, rc, n0, n1 = (int(x) for x in w[i:i + 6])
        blocks = w[i + 6:i + 6 + 2 * (n0 + n1)].reshape(-1, 2)
        reads = [blocks[:n0], blocks[n0:n0 + n1]][:rc]
        got.append("%d\t%d\t%d\t%s" % (chr_, d, rc, "\t".join(",".join("%d-%d" % (a, b) for a, b in r) for r in reads)))
        i += n
    with open(tiny["oracle_frags"]) as f:
        want = [l.rstrip("\n") for l in f]
    assert sorted(got) == sorted(want)


def test_integer_tables_bit_exact(tiny, built, tmp_path):
    c = _core()
    c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
    c.run_bam(tiny["sorted"])
    st = c.stats()
    o = _stats_from_stdout(tiny["oracle_stdout"])
    assert st["pairs"] == int(o["Total pairs processed"]) and st["singles"] == int(o["Total singles processed"])
    assert st["nucleotides"] == int(o["Total nucleotides"]) and st["skipped"] == int(o["Skipped reads"]) and st["error"] == int(o["Error reads"])
    # junction table against the oracle's JuncCount file (counted rows)
    j = c.table("junctions")
    want = {}
    with open(os.path.join(tiny["oracle_out"], "IRFinder-JuncCount.txt")) as f:
        next(f)
        for l in f:
            ch, s, e, _, tot, pos, neg = l.rstrip("\n").split("\t")
            if int(tot):
                want[(ch, int(s), int(e))] = (int(neg), int(pos))
    names = ["chrA", "chrB", "chrEmpty"]
    got = {(names[r["refid"]], int(r["start"]), int(r["end"])): (int(r["neg"]), int(r["pos"])) for r in j}
    assert got == want
    # per-base depth (CoverageBlocks counters after the scan), element for element against the oracle's dump, and through the
    # digest the full-size parity runs use (tools/full_parity.py)
    d = c.table("depth")
    want_d = np.fromfile(tiny["oracle_depth"], dtype="<i4")
    assert d.shape == want_d.shape and d.max() > 0
    assert np.array_equal(d, want_d)
    from conftest import depth_digest
    assert list(depth_digest(d)) == [int(x) for x in open(tiny["oracle_depth_digest"]).read().split()]


def test_stranded_library_writes_dir_table(built, tmp_path):
    meta = _gen(tmp_path, "s", genome="tiny", introns=300, pairs=40000, readlen=100, seed=12, stranded=-1, chunks=5)
    out = run_oracle(built, meta, str(tmp_path / "o"))
    assert "Directionality:\t-1" in out
    c = _gpu_run(meta, str(tmp_path / "g"))
    assert c.directionality == -1 and os.path.exists(str(tmp_path / "g" / "IRFinder-IR-dir.txt"))
    assert_same_outputs(str(tmp_path / "o"), str(tmp_path / "g"))
    meta2 = _gen(tmp_path, "s2", genome="tiny", introns=300, pairs=40000, readlen=100, seed=13, stranded=1, chunks=5)
    out2 = run_oracle(built, meta2, str(tmp_path / "o2"))
    assert "Directionality:\t1" in out2
    _gpu_run(meta2, str(tmp_path / "g2"))
    assert_same_outputs(str(tmp_path / "o2"), str(tmp_path / "g2"))


def test_single_end_dense_introns(built, tmp_path):
    meta = _gen(tmp_path, "se", genome="tiny", introns=200, pairs=60000, readlen=75, seed=14, single=True, intronic=0.4, hot_introns=50, chunks=3)
    run_oracle(built, meta, str(tmp_path / "o"))
    _gpu_run(meta, str(tmp_path / "g"))
    assert_same_outputs(str(tmp_path / "o"), str(tmp_path / "g"))


def test_trim_band_option_and_no_roi(built, tiny, tmp_path):
    run_oracle(built, tiny, str(tmp_path / "o"), env={"IRF_TRIM_CENTER": "60"}, roi=False)
    _gpu_run(tiny, str(tmp_path / "g"), roi=False, trim_center=60)
    assert_same_outputs(str(tmp_path / "o"), str(tmp_path / "g"))
    assert not os.path.exists(str(tmp_path / "g" / "IRFinder-ROI.txt"))


def test_record_aligned_blocks_and_inflated_stream(built, tmp_path):
    meta = _gen(tmp_path, "al", genome="tiny", introns=150, pairs=8000, readlen=100, seed=15, align_records=True, chunks=2)
    run_oracle(built, meta, str(tmp_path / "o"))
    _gpu_run(meta, str(tmp_path / "g"))
    assert_same_outputs(str(tmp_path / "o"), str(tmp_path / "g"))
    # reference-style stdin mode: `gzip -cd < bam | irfinder ... /dev/stdin`
    import gzip
    raw = str(tmp_path / "inflated.bam")
    with gzip.open(meta["sorted"], "rb") as f, open(raw, "wb") as o:
        o.write(f.read())
    c = _core()
    c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
    fd = os.open(raw, os.O_RDONLY)
    try:
        c.run_stream(fd, True)
    finally:
        os.close(fd)
    c.write_outputs(str(tmp_path / "g2"))
    assert_same_outputs(str(tmp_path / "o"), str(tmp_path / "g2"))


def test_cli_shim_matches_reference_invocation(built, tiny, tmp_path):
    exe = os.path.join(ROOT, "bin", "irfinder")
    out = str(tmp_path / "cli")
    os.makedirs(out)
    r = subprocess.run([exe, out, tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"], tiny["sorted"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr
    assert_same_outputs(tiny["oracle_out"], out)
    assert _stats_from_stdout(r.stdout) == _stats_from_stdout(tiny["oracle_stdout"])
    bad = subprocess.run([exe, out], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert bad.returncode != 0


def test_repeated_runs_are_idempotent(tiny, tmp_path):
    c = _core()
    c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
    c.run_bam(tiny["sorted"])
    a = c.table("rows_nondir").tobytes(), c.table("spans").tobytes(), c.table("junctions").tobytes()
    c.run_bam(tiny["collated"])
    b = c.table("rows_nondir").tobytes(), c.table("spans").tobytes(), c.table("junctions").tobytes()
    assert a == b


def test_bad_inputs_fail_loudly(tiny, tmp_path):
    from irfinder_b200 import IRFError
    data = open(tiny["sorted"], "rb").read()
    # de_percent=0: malformed DEFLATE must go to the SM kernel, which reports it and carries on; the hardware decompression
    # engine answers a malformed stream with a sticky CUDA error that would take this whole test process with it
    c = _core(de_percent=0)
    c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
    with pytest.raises(IRFError):   # cut mid-block
        c.run_bytes(data[: len(data) // 2])
    with pytest.raises(IRFError):   # EOF marker missing
        c.run_bytes(data[:-28])
    corrupt = bytearray(data)
    for k in range(40000, 40064):
        corrupt[k] ^= 0x5A           # damage DEFLATE payload
    with pytest.raises(IRFError):
        c.run_bytes(bytes(corrupt))
    with pytest.raises(IRFError):
        c.run_bytes(b"this is not a BAM file at all, not even gzip")
    bad_crc = bytearray(data)     # a block whose payload is intact but whose CRC-32 footer is wrong (what `gzip -cd` would reject)
    bs = (data[16] | data[17] << 8) + 1
    bad_crc[bs - 8] ^= 0x01
    with pytest.raises(IRFError) as ei:
        c.run_bytes(bytes(bad_crc))
    assert "CRC" in str(ei.value)
    c.set_option("verify_crc", 0)
    c.run_bytes(bytes(bad_crc))     # the check can be switched off
    c.set_option("verify_crc", 1)
    c.run_bytes(data)               # the context stays usable afterwards
    assert c.stats()["pairs"] > 0


def _bgzf_block_offsets(data):
    offs, off = [], 0
    while off + 18 <= len(data):
        offs.append(off)
        off += (data[off + 16] | data[off + 17] << 8) + 1
    return offs


@pytest.mark.parametrize("path", ["staged", "pipelined", "pipelined_sm"])
def test_bad_crc_in_a_late_block_on_a_fresh_context(built, tmp_path, path):
    """The integrity flag must survive the table reset that ChrMapUpdate does in the middle of a context's FIRST count (the CLI
    case): a wrong CRC-32 footer in a late block of a multi-MB file is reported on every path.  The damaged block's DEFLATE
    payload is intact, so the hardware decompression engine can take it."""
    from irfinder_b200 import IRFError
    meta = _gen(tmp_path, "w", genome="tiny", introns=300, pairs=120000, readlen=100, seed=5, chunks=7)
    data = bytearray(open(meta["sorted"], "rb").read())
    offs = _bgzf_block_offsets(data)
    assert len(offs) > 100 and len(data) > (6 << 20)
    victim = offs[len(offs) * 9 // 10]
    nxt = offs[len(offs) * 9 // 10 + 1]
    data[nxt - 8] ^= 0x40                      # CRC-32 footer of a block near the end of the file
    opts = dict(staged=dict(stream_pipeline=0), pipelined=dict(stream_chunk_kb=512), pipelined_sm=dict(stream_chunk_kb=512, de_percent=0))[path]
    c = _core(**opts)                          # fresh context: the header is parsed and the references bound inside this very count
    c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
    with pytest.raises(IRFError) as ei:
        c.run_bytes(bytes(data))
    assert "CRC" in str(ei.value), (path, victim, str(ei.value))
    data[nxt - 8] ^= 0x40
    c.run_bytes(bytes(data))                   # and the context recovers
    assert c.stats()["pairs"] > 0


def test_cli_default_options_on_damaged_bams(built, tiny, tmp_path):
    """"Truncated BGZF / bad CRC must be a hard error with non-zero exit" — through bin/irfinder with DEFAULT options
    (every block goes to the hardware decompression engine).  Each damaged file runs in its own process: non-zero exit, a message
    on stderr, no hang; a fresh process afterwards works, and IRF_SAFE_INFLATE=1 names the problem without a CUDA fault."""
    exe = os.path.join(ROOT, "bin", "irfinder")
    data = open(tiny["sorted"], "rb").read()
    offs = _bgzf_block_offsets(data)
    mid = offs[len(offs) // 2]
    cases = {}
    cases["truncated_mid_block"] = data[: mid + 1000]
    flipped = bytearray(data)
    for k in range(mid + 400, mid + 464):
        flipped[k] ^= 0xA5                      # DEFLATE payload of a block in the middle of the file
    cases["bit_flipped_payload"] = bytes(flipped)
    one_bit = bytearray(data)
    one_bit[mid + 700] ^= 0x10
    cases["single_bit_flip"] = bytes(one_bit)
    args = [tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"]]
    for name, blob in cases.items():
        bam = str(tmp_path / (name + ".bam"))
        open(bam, "wb").write(blob)
        for env_extra in ({}, {"IRF_SAFE_INFLATE": "1"}):
            out = str(tmp_path / (name + ("_safe" if env_extra else "")))
            os.makedirs(out, exist_ok=True)
            r = subprocess.run([exe, out] + args + [bam], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=60, env=dict(os.environ, **env_extra))
            assert r.returncode != 0, (name, env_extra, r.stdout[-300:])
            assert r.stderr.strip(), (name, env_extra)
            if env_extra:
                assert "CUDA" not in r.stderr and "cuda" not in r.stderr, (name, r.stderr)    # the SM kernel reports a format error, not a device fault
    out = str(tmp_path / "good")
    os.makedirs(out)
    r = subprocess.run([exe, out] + args + [tiny["sorted"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    assert_same_outputs(tiny["oracle_out"], out)


@pytest.mark.parametrize("order", ["sorted", "collated"])
@pytest.mark.parametrize("inflated", [0, 1])
def test_stream_input_is_counted_segment_by_segment(built, tiny, tmp_path, order, inflated):
    """irf_run_stream (the reference's /dev/stdin mode) reads a pipe of unknown length in bounded segments:
    here 1 MB segments of a 7 MB file, so records straddle segment ends and (in the coordinate-sorted file) mates sit in different
    segments and are carried over.  BGZF bytes and an already-inflated stream, both record orders; a stream cut short must fail."""
    from bamutil import inflate_file
    from irfinder_b200 import IRFError
    path = tiny[order]
    if inflated:
        path = str(tmp_path / "raw.bam.bytes")
        open(path, "wb").write(inflate_file(tiny[order]))
    c = _core(stream_segment_mb=1)
    c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
    for rep in range(2):                                   # twice on one context: nothing of the first run may leak into the second
        r, w = os.pipe()
        feeder = subprocess.Popen(["dd", "if=" + path, "bs=777777", "status=none"], stdout=w)   # a real pipe, fed in odd-sized pieces
        os.close(w)
        c.run_stream(r, inflated)
        os.close(r)
        assert feeder.wait() == 0
        out = str(tmp_path / f"g{rep}")
        c.write_outputs(out)
        assert_same_outputs(tiny["oracle_out"], out)
        assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(tiny["oracle_stdout"])
    cut = str(tmp_path / "cut.bytes")
    data = open(path, "rb").read()
    open(cut, "wb").write(data[: len(data) * 2 // 3 + 11])
    fd = os.open(cut, os.O_RDONLY)
    with pytest.raises(IRFError):
        c.run_stream(fd, inflated)
    os.close(fd)
    fd = os.open(path, os.O_RDONLY)
    c.run_stream(fd, inflated)                             # the context stays usable
    os.close(fd)
    assert c.stats()["pairs"] > 0


_RULE_SETS = {
    "junction_span_flags": dict(junc_min_flank=4, span_left=4, span_right=5, skip_flags=0x900),
    "percentile_edge": dict(pct_tie_midpoint=0, edge_offset=3, edge_len=40),
    "warnings": dict(warn_lowcover_e6=5_000_000, warn_lowsplice_e6=2_000_000, warn_minor_ratio_e6=250_000, warn_nonuniform_abs_e6=2_000_000, warn_nonuniform_ratio_e6=1_200_000),
}


@pytest.mark.parametrize("which", sorted(_RULE_SETS))
def test_appendix_c_rules_are_run_time_settings_on_both_sides(built, tiny, tmp_path, which):
    """SURVEY.md Appendix C / §8c "Mitigation": every behaviour the recollected specification is unsure about (junction flank, span
    overhangs, secondary/supplementary handling, percentile tie rule, edge windows, warning thresholds) is a run-time setting in the
    library (irf_set_option rule_*) and, independently, in the oracle (IRF_RULE_*).  A second setting of each must agree byte for
    byte too, and must actually change the output (otherwise the knob is not wired)."""
    rules = _RULE_SETS[which]
    env = {"IRF_RULE_" + k.upper(): str(v) for k, v in rules.items()}
    out_o = str(tmp_path / "o")
    ostd = run_oracle(built, tiny, out_o, env=env)
    c = _gpu_run(tiny, str(tmp_path / "g"), **{"rule_" + k: v for k, v in rules.items()})
    assert_same_outputs(out_o, str(tmp_path / "g"))
    assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(ostd)
    changed = [n for n in ("IRFinder-IR-nondir.txt", "IRFinder-JuncCount.txt", "IRFinder-SpansPoint.txt")
               if open(os.path.join(out_o, n), "rb").read() != open(os.path.join(tiny["oracle_out"], n), "rb").read()]
    assert changed, "the alternative rules left every table unchanged"
    c.set_option("rule_junc_min_flank", 5); c.set_option("rule_span_left", 5); c.set_option("rule_span_right", 4); c.set_option("rule_skip_flags", 0)
    c.set_option("rule_pct_tie_midpoint", 1); c.set_option("rule_edge_offset", 5); c.set_option("rule_edge_len", 50)
    for k, v in dict(warn_lowcover_e6=10_000_000, warn_lowsplice_e6=4_000_000, warn_minor_ratio_e6=333_330, warn_nonuniform_abs_e6=4_000_000, warn_nonuniform_ratio_e6=1_500_000).items():
        c.set_option("rule_" + k, v)
    c.run_bam(tiny["sorted"])                       # back to the defaults on the same context: the default tables again
    c.write_outputs(str(tmp_path / "d"))
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "d"))


@pytest.mark.parametrize("mode", [1, 2])
def test_depth_tile_modes_are_bit_exact(built, tiny, tmp_path, mode):
    """CoverageBlocks counters summed in a shared-memory tile per CTA before they reach the difference table (option depth_tile: 1 = tile
    flushed with one red per non-zero counter, 2 = flushed by the TMA unit, cp.reduce.async.bulk add.s32): the per-base depth must be
    bit-identical to the oracle's on a paired and on a dense single-end workload, through the staged path and the chunk pipeline."""
    c = _core(depth_tile=mode)
    c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
    for bam in (tiny["sorted"], tiny["collated"]):
        c.run_bam(bam)
        assert np.array_equal(c.table("depth"), np.fromfile(tiny["oracle_depth"], dtype="<i4"))
    c.write_outputs(str(tmp_path / "g"))
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "g"))
    meta = _gen(tmp_path, "se", genome="tiny", introns=200, pairs=60000, readlen=75, seed=14, single=True, intronic=0.4, hot_introns=50, chunks=3)
    run_oracle(built, meta, str(tmp_path / "o2"), depth=str(tmp_path / "d2.i32"))
    c2 = _gpu_run(meta, str(tmp_path / "g2"), depth_tile=mode, stream_chunk_kb=256)
    assert_same_outputs(str(tmp_path / "o2"), str(tmp_path / "g2"))
    assert np.array_equal(c2.table("depth"), np.fromfile(str(tmp_path / "d2.i32"), dtype="<i4"))


def test_micro_fixture_by_hand_on_gpu(tmp_path):
    """The hand-derived fixture of tests/micro_fixture.py (every cell computed by hand) through the CUDA path,
    in both record orders; BGZF blocks of 700-900 bytes make nearly every record straddle a block boundary."""
    import micro_fixture
    p = micro_fixture.write(str(tmp_path / "m"))
    for key in ("collated", "sorted"):
        c = _core()
        c.load_ref(p["cover"], p["sj"], p["rc"], p["roi"])
        c.run_bam(p[key])
        out = str(tmp_path / ("g_" + key))
        c.write_outputs(out)
        for name, want in micro_fixture.EXPECT.items():
            assert open(os.path.join(out, name)).read() == want, (key, name)
        st = _stats_from_stdout(c.stdout_text)
        for k, v in micro_fixture.EXPECT_STATS.items():
            assert st[k] == v, (key, k)


def test_rule_probe_fixture_by_hand_on_gpu(tmp_path):
    """tests/micro_fixture2.py through the CUDA path: the default table cell for cell as derived by hand, then every alternative rule
    (irf_set_option rule_*) changes exactly the hand-derived cells.  Independent of the oracle: the expectations are literals."""
    import micro_fixture2
    from test_oracle_cpu import _check_rule_probe
    p = micro_fixture2.write(str(tmp_path / "m"))
    for key in ("collated", "sorted"):
        c = _core()
        c.load_ref(p["cover"], p["sj"], p["rc"], p["roi"])
        c.run_bam(p[key])
        out = str(tmp_path / ("g_" + key))
        c.write_outputs(out)
        for name, want in micro_fixture2.EXPECT.items():
            assert open(os.path.join(out, name)).read() == want, (key, name)
        st = _stats_from_stdout(c.stdout_text)
        for k, v in micro_fixture2.EXPECT_STATS.items():
            assert st[k] == v, (key, k)
    for which, (rules, _) in micro_fixture2.RULE_PROBES.items():
        c = _core(**{"rule_" + k: v for k, v in rules.items()})
        c.load_ref(p["cover"], p["sj"], p["rc"], p["roi"])
        c.run_bam(p["sorted"])
        out = str(tmp_path / ("r_" + which))
        c.write_outputs(out)
        _check_rule_probe(open(os.path.join(out, "IRFinder-IR-nondir.txt")).read(), which)


def test_golden_vectors_on_gpu(built, tmp_path):
    meta = _gen(tmp_path, "w", genome="tiny", introns=120, pairs=4000, readlen=100, seed=21, chunks=3)
    _gpu_run(meta, str(tmp_path / "g"))
    assert_same_outputs(os.path.join(ROOT, "tests", "golden", "tiny_seed21"), str(tmp_path / "g"))


@pytest.mark.parametrize("world", [2, 3, 5, "3+bai"])
def test_region_sharded_run_is_byte_identical(tiny, tmp_path, world):
    """shard by BGZF byte range, frame hand-shake, halo exchange of cross-shard mates, junction-table and
    counter merge — emulated with `world` contexts on one GPU — must reproduce the unsharded outputs byte for byte.
    "3+bai": the cut points are read from a BAI index of the file instead of being found by scanning for block headers."""
    from irfinder_b200.distributed import run_sharded_local
    bai = None
    if world == "3+bai":
        from bamutil import write_bai
        world, bai = 3, write_bai(tiny["sorted"], str(tmp_path / "sorted.bam.bai"))
    cores = []
    for _ in range(world):
        c = _core()
        c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
        cores.append(c)
    buf = np.fromfile(tiny["sorted"], dtype=np.uint8)
    info = run_sharded_local(cores, buf, bai=bai)
    if bai is None:      # index cuts sit where a 16 kb window begins, which in this small file can fall between read clusters
        assert info["halo_records"] > 0, "the cut points must split at least one pair for this test to mean anything"
    if world == 5 and bai is None:   # region-owned depth: with five shards of this small genome some coverage must land in another rank's segments
        assert info["foreign_depth_entries"] > 0, "no counter crossed an ownership boundary: the sparse depth exchange was not exercised"
    for r, c in enumerate(cores):
        out = str(tmp_path / f"g{r}")
        c.write_outputs(out)
        assert_same_outputs(tiny["oracle_out"], out)
        assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(tiny["oracle_stdout"])


def test_wrapper_script_bam_mode(built, tiny, tmp_path):
    """`IRFinder -m BAM -r <RefDir> -d <OutDir> <bam>`."""
    exe = os.path.join(ROOT, "bin", "IRFinder")
    out = str(tmp_path / "wrapped")
    r = subprocess.run([exe, "-m", "BAM", "-r", tiny["ref_dir"], "-d", out, tiny["sorted"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stderr + open(os.path.join(out, "irfinder.stderr")).read()
    assert_same_outputs(tiny["oracle_out"], out)
    assert _stats_from_stdout(open(os.path.join(out, "irfinder.stdout")).read()) == _stats_from_stdout(tiny["oracle_stdout"])
    again = subprocess.run([exe, "-m", "BAM", "-r", tiny["ref_dir"], "-d", out, tiny["sorted"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert again.returncode != 0      # refuses to overwrite
    other = subprocess.run([exe, "-m", "FastQ", "-r", tiny["ref_dir"], "-d", str(tmp_path / "x"), tiny["sorted"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
    assert other.returncode != 0


def test_baseline_config0_chr22_1m_pairs(built, tmp_path):
    """BASELINE.json configs[0]: 1M paired-end 2x100 over a 50 Mb chr22-like genome (~5k introns) — the reference-runnable
    case — GPU on the coordinate-sorted BAM vs the oracle on the name-collated copy, every output file byte for byte."""
    meta = _gen(tmp_path, "c1", genome="chr22", introns=5000, pairs=1_000_000, readlen=100, seed=1, chunks=64)
    out = run_oracle(built, meta, str(tmp_path / "o"))
    c = _gpu_run(meta, str(tmp_path / "g"))
    assert_same_outputs(str(tmp_path / "o"), str(tmp_path / "g"))
    assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(out)


def test_deep_introns_use_the_global_histogram_path(built, tmp_path):
    """Depths beyond the 8192-bin shared-memory histogram (BASELINE.json configs[3]: dense intronic coverage) take the
    per-CTA global scratch histogram in intron_stats_kernel; statistics must still match the oracle exactly."""
    import micro_fixture
    from bamutil import record, write_bam
    p = micro_fixture.write(str(tmp_path / "m"))
    recs = [record("d%05d" % i, 0, 1400 + (i % 300), "100M", 16 * (i & 1)) for i in range(12000)]          # ~4000x over intron 1
    recs += [record("e%05d" % i, 0, 3100 + (i % 7), "120M", 0) for i in range(9500)]                       # ~9500x over intron 2
    recs += [record("f%05d" % i, 0, 900, "100M1000N50M", 0) for i in range(40)]
    bam = str(tmp_path / "deep.bam")
    write_bam(bam, micro_fixture.REFS, recs, level=1)
    meta = dict(cover=p["cover"], sj=p["sj"], rc=p["rc"], roi=p["roi"], collated=bam, sorted=bam)
    run_oracle(built, meta, str(tmp_path / "o"))
    c = _gpu_run(meta, str(tmp_path / "g"))
    d = c.table("depth"); half = d.size // 2
    assert int((d[:half] + d[half:]).max()) >= 8192, "workload too shallow to exercise the scratch path"
    assert_same_outputs(str(tmp_path / "o"), str(tmp_path / "g"))


@pytest.mark.parametrize("de_percent", [0, 35, 100])
def test_inflate_engines_agree(tiny, tmp_path, de_percent):
    """BGZF blocks may be inflated by the SM kernel, by the Blackwell hardware decompression engine, or split between them:
    the outputs must not depend on the split (and CRC-32 of every block is verified in all cases)."""
    c = _gpu_run(tiny, str(tmp_path / "g"), de_percent=de_percent)
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "g"))
    assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(tiny["oracle_stdout"])


@pytest.mark.parametrize("chunk_kb", [64, 300])
def test_streaming_pipeline_counts_chunk_by_chunk(tiny, tmp_path, chunk_kb):
    """Host-resident input is copied, inflated, framed and counted chunk by chunk (count_streaming).  Tiny chunks put a chunk
    boundary after every few BGZF blocks, so records and mate pairs straddle chunks all the time; outputs must not change."""
    c = _core(stream_chunk_kb=chunk_kb)
    c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
    for bam in (tiny["sorted"], tiny["collated"]):
        c.run_bam(bam)
        out = str(tmp_path / ("g_" + os.path.basename(bam)))
        c.write_outputs(out)
        assert_same_outputs(tiny["oracle_out"], out)
        assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(tiny["oracle_stdout"])
    c.set_option("stream_pipeline", 0)     # the unpipelined path must agree
    c.run_bam(tiny["sorted"])
    c.write_outputs(str(tmp_path / "g_plain"))
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "g_plain"))


def test_malformed_reference_files_are_reported_with_line_numbers(tiny, tmp_path):
    from irfinder_b200 import IRFError
    c = _core()
    lines = open(tiny["cover"]).read().split("\n")
    for mutate, needle in ((lambda f: f[:11], "12 BED columns"), (lambda f: f[:9] + ["7"] + f[10:], "blockCount"), (lambda f: f[:1] + ["12x"] + f[2:], "bad start")):
        f = lines[2].split("\t")
        bad = str(tmp_path / "bad-cover.bed")
        open(bad, "w").write("\n".join(lines[:2] + ["\t".join(mutate(f))] + lines[3:]))
        with pytest.raises(IRFError) as ei:
            c.load_ref(bad, tiny["sj"], tiny["rc"], tiny["roi"])
        assert "ref-cover.bed line 3" in str(ei.value) and needle in str(ei.value)
    with pytest.raises(IRFError):
        c.load_ref(tiny["cover"], str(tmp_path / "missing.ref"), tiny["rc"], tiny["roi"])
    c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])   # a good set still loads afterwards
    c.run_bam(tiny["sorted"])
    assert c.stats()["pairs"] > 0


@pytest.mark.parametrize("share", [30, 65, 100])
def test_resident_pipeline_overlaps_engines(tiny, tmp_path, share):
    """Bytes already on the GPU are re-counted by the overlapped pipeline (count_streaming(false)): the SM kernel inflates the
    leading chunks, the decompression engine the rest chunk by chunk, counting follows both.  Small chunks make the tiny input
    span enough chunks to take that path; the share of the two engines and repeated counts must not change any output."""
    import numpy as np
    c = _core(stream_chunk_kb=64, de_percent_overlap=share, resident_pipeline=2)
    c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
    data = np.fromfile(tiny["sorted"], dtype=np.uint8)
    assert data.size >= 8 * 65536          # otherwise the path under test is not taken
    c.stage(data.ctypes.data, data.size)
    for rep in range(2):
        c.reset(); c.count(); c.finalize()
        out = str(tmp_path / ("g%d" % rep))
        c.write_outputs(out)
        assert_same_outputs(tiny["oracle_out"], out)
        assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(tiny["oracle_stdout"])
        assert c.timing()["inflate_launches"] == (0 if share == 100 else 1)
    c.set_option("resident_pipeline", 0)
    c.reset(); c.count(); c.finalize()
    c.write_outputs(str(tmp_path / "plain"))
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "plain"))


@pytest.mark.parametrize("mode", ["resident", "streamed"])
def test_sharded_run_through_the_chunk_pipeline(tiny, tmp_path, mode):
    """Shards (no header, look-ahead blocks, guessed first record) counted by the chunk pipeline — overlapped two-engine
    inflate for resident bytes, copy/inflate/count for bytes streamed from the host — must match the unsharded outputs."""
    from irfinder_b200.distributed import run_sharded_local
    cores = []
    for _ in range(3):
        opts = dict(stream_chunk_kb=64, resident_pipeline=2, de_percent_overlap=60) if mode == "resident" else dict(stream_chunk_kb=64, stage_lazy=1)
        c = _core(**opts)
        c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
        cores.append(c)
    buf = np.fromfile(tiny["sorted"], dtype=np.uint8)
    info = run_sharded_local(cores, buf)
    assert info["halo_records"] > 0
    for r, c in enumerate(cores):
        assert c.timing()["framing_ms"] == 0.0, "the chunk pipeline reports no separate framing stage: the stage-by-stage path ran instead"
        out = str(tmp_path / f"g{r}")
        c.write_outputs(out)
        assert_same_outputs(tiny["oracle_out"], out)
        assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(tiny["oracle_stdout"])


def _adversarial_records(sj_path, seed):
    """Hand-made alignments that a read simulator never produces, around the annotated junctions of the tiny genome."""
    import random
    from bamutil import record
    rnd = random.Random(seed)
    junc = [l.split("\t") for l in open(sj_path).read().split("\n") if l]
    junc = [(0 if j[0] == "chrA" else 1, int(j[1]), int(j[2])) for j in junc if j[0] in ("chrA", "chrB")]
    P, PROPER, UNMAP, MUNMAP, REV, MREV, R1, R2, SEC, QCF, DUP, SUP = 0x1, 0x2, 0x4, 0x8, 0x10, 0x20, 0x40, 0x80, 0x100, 0x200, 0x400, 0x800
    frags = []          # one entry per fragment: mates stay next to each other in the order the oracle reads

    class _Sink:
        def append(self, r):
            frags.append([r])
    recs = _Sink()

    def pair(name, refid, p1, c1, p2, c2, f1=0, f2=0, mref=None):
        a = record(name, refid, p1, c1, P | PROPER | R1 | f1, refid if mref is None else mref, p2, 300)
        b = record(name, refid if mref is None else mref, p2, c2, P | PROPER | R2 | f2, refid, p1, -300)
        frags.append([a, b] if rnd.randrange(2) else [b, a])

    for i in range(1500):
        ref, s, e = junc[rnd.randrange(len(junc))]
        gap = e - s
        kind = rnd.randrange(14)
        a, b = rnd.randrange(1, 60), rnd.randrange(1, 60)            # flanks below and above the 5-base threshold
        name = "q%05d" % i
        if kind == 0:    # spliced read exactly on an annotated junction, short and long flanks, both strands
            recs.append(record(name, ref, s - a, f"{a}M{gap}N{b}M", REV * rnd.randrange(2)))
        elif kind == 1:  # every CIGAR operation: clips, insertion, deletion, padding, = and X, two gaps
            recs.append(record(name, ref, s - 30, f"3H4S10=2X1I8M2D10M{gap}N12M1P5M40N9M6S", REV * rnd.randrange(2)))
        elif kind == 2:  # pair whose reads overlap, read 2 first
            pair(name, ref, s - 80, "100M", s - 40, "100M", MREV, REV)
        elif kind == 3:  # read 2 contained in read 1, both spliced over the same junction
            pair(name, ref, s - 50, f"50M{gap}N50M", s - 20, f"20M{gap}N10M", MREV, REV)
        elif kind == 4:  # far apart mates (no overlap), reverse read first in the file
            pair(name, ref, max(0, s - 900), "75M", s - 10, f"10M{gap}N65M", REV, MREV)
        elif kind == 5:  # mates on different chromosomes
            pair(name, ref, s - 10, "60M", 1000 + i, "60M", mref=2)
        elif kind == 6:  # mate flagged unmapped (placed at its mate's position, as aligners do)
            recs.append(record(name, ref, s - 20, "70M", P | MUNMAP | R1))
            recs.append(record(name, ref, s - 20, "", P | UNMAP | R2, lseq=70))
        elif kind == 7:  # secondary / supplementary / QC-fail / duplicate flags on otherwise ordinary reads
            recs.append(record(name, ref, s - 30, f"30M{gap}N40M", rnd.choice([SEC, SUP, QCF, DUP, SEC | REV])))
        elif kind == 8:  # the same read name used by two different pairs
            pair("shared", ref, s - 60, "50M", e + 5, "50M", MREV, REV)
        elif kind == 9:  # a paired flag with the mate never present in the file
            recs.append(record(name, ref, e - 10, "90M", P | PROPER | R1 | MREV, ref, e + 200, 300))
        elif kind == 10:  # read ending exactly at / starting exactly at the splice sites
            recs.append(record(name, ref, s - 50, "50M", 0)); recs.append(record(name + "b", ref, e, "50M", REV))
        elif kind == 11:  # novel junction sharing one annotated end
            recs.append(record(name, ref, s - 25, f"25M{gap + 7}N25M", 0))
        elif kind == 12:  # intron-body reads at both edges of the measured region
            recs.append(record(name, ref, s + rnd.randrange(0, 8), "50M", 0)); recs.append(record(name + "b", ref, max(s, e - 50 - rnd.randrange(0, 8)), "50M", REV))
        else:             # deletion instead of a gap across the whole intron (counts as coverage, not as a junction)
            recs.append(record(name, ref, s - 10, f"10M{min(gap, 2000)}D10M", 0))
    recs.append(record("first_base", 0, 0, "80M", 0))
    recs.append(record("last_base", 1, 300000 - 80, "80M", REV))
    recs.append(record("empty_chr", 2, 100, "80M", 0))
    recs.append(record("long_read", 0, 50000, "60000M5000N90000M", 0))             # one 225 KB record across four BGZF blocks
    recs.append(record("unmapped_tail", -1, -1, "", UNMAP, lseq=100))
    rnd.shuffle(frags)
    collated = [r for f in frags for r in f]
    scattered = list(collated)
    rnd.shuffle(scattered)          # mates far apart, in any order: only the GPU path (hash pairing) reads this one
    return collated, scattered


@pytest.mark.parametrize("block", [65280, 3000])
def test_adversarial_alignments_match_oracle(built, tiny, tmp_path, block):
    """Flag combinations, CIGAR operations, name collisions, cross-chromosome and orphan mates, reads at chromosome ends and a
    read longer than three BGZF blocks, shuffled, in normal and in very small BGZF blocks: GPU tables must equal the oracle's."""
    from bamutil import write_bam
    refs = [("chrA", 400000), ("chrB", 300000), ("chrEmpty", 50000)]
    collated, scattered = _adversarial_records(tiny["sj"], 5)
    bam, bam2 = str(tmp_path / "adv.bam"), str(tmp_path / "adv_scattered.bam")
    write_bam(bam, refs, collated, block=block, level=1)
    write_bam(bam2, refs, scattered, block=block, level=1)
    meta = dict(tiny, sorted=bam, collated=bam)
    ostd = run_oracle(built, meta, str(tmp_path / "o"))
    want = _stats_from_stdout(ostd)
    assert int(want["Total pairs processed"]) > 300 and int(want["Total singles processed"]) > 500 and int(want["Skipped reads"]) > 50 and int(want["Error reads"]) > 50
    for i, (path, opts) in enumerate([(bam, {}), (bam2, {}), (bam2, dict(stream_chunk_kb=64, de_percent=0, de_percent_stream=0)),
                                      (bam2, dict(stream_chunk_kb=64))]):      # chunk pipeline: the long read spans several 64 KB chunks
        c = _gpu_run(meta, str(tmp_path / f"g{i}"), bam=path, **opts)
        assert_same_outputs(str(tmp_path / "o"), str(tmp_path / f"g{i}"))
        assert _stats_from_stdout(c.stdout_text) == want


def test_header_only_and_many_contig_inputs(built, tiny, tmp_path):
    """Ragged inputs: a BAM with a header and no alignments at all, and one whose header (4000 contigs, > 64 KB) spans several
    BGZF blocks with alignments only on the annotated chromosomes and on the last contig."""
    from bamutil import record, write_bam
    refs = [("chrA", 400000), ("chrB", 300000), ("chrEmpty", 50000)]
    empty = str(tmp_path / "empty.bam")
    write_bam(empty, refs, [], level=1)
    meta = dict(tiny, sorted=empty, collated=empty)
    ostd = run_oracle(built, meta, str(tmp_path / "o0"))
    c = _gpu_run(meta, str(tmp_path / "g0"))
    assert_same_outputs(str(tmp_path / "o0"), str(tmp_path / "g0"))
    assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(ostd)
    assert c.stats()["records"] == 0

    many = refs + [("scaffold_%04d_random_with_a_long_name" % i, 1000 + i) for i in range(4000)]
    junc = [l.split("\t") for l in open(tiny["sj"]).read().split("\n") if l][:200]
    recs = []
    for i, j in enumerate(junc):
        ref, s, e = (0 if j[0] == "chrA" else 1), int(j[1]), int(j[2])
        recs.append(record("m%04d" % i, ref, s - 40, f"40M{e - s}N40M", 16 * (i & 1)))
        recs.append(record("s%04d" % i, 3 + (i * 19) % 4000, 10, "50M", 0))
    recs.append(record("last", len(many) - 1, 100, "30M", 16))
    big = str(tmp_path / "many.bam")
    raw = write_bam(big, many, recs, level=1)
    assert raw.index(b"m0000\0") > 2 * 65280, "the header must span more than two BGZF blocks for this test to mean anything"
    meta = dict(tiny, sorted=big, collated=big)
    ostd = run_oracle(built, meta, str(tmp_path / "o1"))
    for i, opts in enumerate([{}, dict(stream_chunk_kb=64), dict(de_percent=0, de_percent_stream=0)]):
        c = _gpu_run(meta, str(tmp_path / f"g1_{i}"), **opts)
        assert_same_outputs(str(tmp_path / "o1"), str(tmp_path / f"g1_{i}"))
        assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(ostd)


def test_decoy_record_at_a_block_start_is_repaired_in_one_round(built, tiny, tmp_path):
    """Record framing guesses each BGZF block's first record from the byte layout.  Here the quality string of a long read
    carries, exactly at a block start, bytes that look like two well-formed records whose chain leaves the block at a wrong
    offset.  The guess for that block is therefore wrong; the repair must fix it from the predecessor's exit without
    overwriting the (correct) guesses of the 300+ blocks that follow, i.e. in a couple of rounds, not one round per block."""
    import struct
    from bamutil import BGZF_EOF, bgzf, header, record
    refs = [("chrA", 400000), ("chrB", 300000), ("chrEmpty", 50000)]
    B = 2048
    head = header(refs, True) + b"".join(record("p%04d" % i, 0, 1000 + i, "100M", 0) for i in range(40))
    long_read = bytearray(record("decoy_carrier", 0, 5000, "6000M", 0))
    qual0 = len(long_read) - 4 - 6000            # quality string = the 6000 bytes before the 4-byte tag
    k = (len(head) + qual0 + B - 1) // B + 1     # a block start well inside the quality string
    at = k * B - len(head)
    assert qual0 < at and at + 200 < qual0 + 6000

    def fake(block_size, name=b"abc\0"):
        return struct.pack("<iiiBBHHHiiii", block_size, 0, 100, len(name), 30, 4680, 1, 0, 10, -1, -1, 0) + name + struct.pack("<I", 10 << 4)
    f1, f2 = fake(60), fake(B + 777)
    long_read[at:at + len(f1)] = f1
    long_read[at + 64:at + 64 + len(f2)] = f2     # f1's chain: next record at +4+60
    tail = b"".join(record("t%05d" % i, 0, 20000 + 3 * i, "60M500N60M" if i % 5 == 0 else "120M", 16 * (i & 1)) for i in range(3500))
    raw = head + bytes(long_read) + tail
    assert len(raw) // B - k > 300
    bam = str(tmp_path / "decoy.bam")
    with open(bam, "wb") as f:
        f.write(bgzf(raw, B, 1) + BGZF_EOF)
    meta = dict(tiny, sorted=bam, collated=bam)
    ostd = run_oracle(built, meta, str(tmp_path / "o"))
    for i, opts in enumerate([{}, dict(stream_pipeline=0), dict(stream_pipeline=0, de_percent=0, de_percent_stream=0)]):
        c = _gpu_run(meta, str(tmp_path / f"g{i}"), **opts)
        assert_same_outputs(str(tmp_path / "o"), str(tmp_path / f"g{i}"))
        assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(ostd)
        assert c.timing()["launches"] < 120, "framing repair ran one round per block"


@pytest.mark.parametrize("chunk_kb", [0, 64])
def test_file_input_through_the_page_locked_ring(tiny, tmp_path, chunk_kb):
    """irf_run_bam on a regular file: the file is mapped for the block scan and its bytes reach the GPU through a small ring of
    page-locked buffers filled by pread threads (forced here for a small file with file_ring=2).  Same outputs; the ring
    is reused from run to run; a file that fails validation while the feeder threads are running leaves the context usable."""
    from irfinder_b200 import IRFError
    opts = dict(file_ring=2)
    if chunk_kb:
        opts["stream_chunk_kb"] = chunk_kb
    c = _core(**opts)
    c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
    for i, bam in enumerate([tiny["sorted"], tiny["collated"], tiny["sorted"]]):
        c.run_bam(bam)
        out = str(tmp_path / f"g{i}")
        c.write_outputs(out)
        assert_same_outputs(tiny["oracle_out"], out)
        assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(tiny["oracle_stdout"])
    cut = str(tmp_path / "cut.bam")
    data = open(tiny["sorted"], "rb").read()
    open(cut, "wb").write(data[: len(data) // 2])
    with pytest.raises(IRFError):
        c.run_bam(cut)
    c.run_bam(tiny["sorted"])
    c.write_outputs(str(tmp_path / "again"))
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "again"))
    c.reset(); c.count(); c.finalize()          # the bytes stay resident after a file run as after any other
    c.write_outputs(str(tmp_path / "recount"))
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "recount"))


@pytest.mark.parametrize("variant", [0, 12, 13, 14, 15, 16, 17, 10, 11, 7, 2])
def test_inflate_kernel_variants_agree(tiny, tmp_path, variant):
    """The A/B variants of the SM inflate kernel (pre-written sectors, wider LUTs, byte-entry distance LUT with five or six
    CTAs per SM; the shared-memory output-ring variants are disabled, see launch_inflate) must produce the same bytes as the default: outputs equal the oracle's with the
    decompression engine switched off, so every block goes through the variant under test, and every CRC-32 is checked."""
    c = _gpu_run(tiny, str(tmp_path / "g"), inflate_variant=variant, de_percent=0, de_percent_stream=0)
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "g"))
    assert c.timing()["inflate_launches"] >= 1 and c.timing()["sm_inflate_share"] == 1.0


def test_default_sm_inflate_is_the_two_phase_kernel(tiny, tmp_path):
    """The default inflate_variant must launch inflate_tok_kernel (a dispatch slip once sent it to the one-phase kernel: same bytes, 25 % slower,
    and no parity test can see that), variant 0 must not, and an unknown variant is refused."""
    from irfinder_b200 import IRFError
    c = _gpu_run(tiny, str(tmp_path / "g"), de_percent=0, de_percent_stream=0)
    before = c.timing()["tok_launches_total"]
    c.run_bam(tiny["sorted"])
    assert c.timing()["tok_launches_total"] > before
    c0 = _gpu_run(tiny, str(tmp_path / "g0"), inflate_variant=0, de_percent=0, de_percent_stream=0)
    mid = c0.timing()["tok_launches_total"]
    c0.run_bam(tiny["sorted"])
    assert c0.timing()["tok_launches_total"] == mid
    bad = _core(inflate_variant=99, de_percent=0, de_percent_stream=0)
    bad.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
    with pytest.raises(IRFError):
        bad.run_bam(tiny["sorted"])


@pytest.mark.parametrize("level,block", [(0, 65280), (6, 65280), (9, 65280), (6, 1500), (1, 20000)])
def test_two_phase_inflate_on_other_encoders(built, tiny, tmp_path, level, block):
    """The two-phase SM kernel (token decode per lane, warp-wide match resolve) on BGZF written by other encoder settings than the
    generator's zlib level 1: stored blocks (level 0: literals only), levels 6 and 9 (long matches, lazy matching, run-length matches of
    distance 1 in the header text and the quality strings), and small blocks (many blocks per warp, output words shared by neighbouring blocks)."""
    from bamutil import BGZF_EOF, bgzf, inflate_file
    raw = inflate_file(tiny["sorted"])
    assert len(raw) > 1 << 20
    bam = str(tmp_path / "re.bam")
    open(bam, "wb").write(bgzf(raw, block=block, level=level) + BGZF_EOF)
    c = _gpu_run(tiny, str(tmp_path / "g"), bam=bam, inflate_variant=12, de_percent=0, de_percent_stream=0)
    assert_same_outputs(tiny["oracle_out"], str(tmp_path / "g"))
    assert c.timing()["sm_inflate_share"] == 1.0


# ---- in-process multi-device context (irf_create with n_devices > 1, csrc/multi.cu; SURVEY.md §8b "one ctx drives all its GPUs") ----

def _device_lists():
    """Several shards on GPU 0 always; real device lists as far as the box has GPUs."""
    import torch
    n = torch.cuda.device_count()
    lists = [[0, 0], [0, 0, 0], [0] * 5]
    if n >= 2:
        lists += [[0, 1], list(range(n))]
    return lists


@pytest.mark.parametrize("ring", [0, 2])
def test_one_context_drives_several_devices(built, tiny, tmp_path, ring):
    """irf_create(device_ids, n > 1): irf_run_bam / irf_run_buffer cut the BAM into one byte range per entry, count them side by side on
    host threads, merge (hand-shake, halo, junctions, counters, region-owned depth, rows) and leave the complete tables in the
    context - byte-identical to the oracle, for files read through the page-locked ring (ring=2) and through the mapping (ring=0)."""
    from irfinder_b200 import IRFinderCore, IRFError
    for k, devs in enumerate(_device_lists()):
        c = IRFinderCore(devs, file_ring=ring, stream_chunk_kb=64)
        c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
        for i, bam in enumerate([tiny["sorted"], tiny["collated"]]):
            c.run_bam(bam)
            out = str(tmp_path / f"g{k}_{i}")
            c.write_outputs(out)
            assert_same_outputs(tiny["oracle_out"], out)
            assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(tiny["oracle_stdout"])
        buf = np.fromfile(tiny["sorted"], dtype=np.uint8)
        c.run_buffer(buf.ctypes.data, buf.size)
        c.write_outputs(str(tmp_path / f"b{k}"))
        assert_same_outputs(tiny["oracle_out"], str(tmp_path / f"b{k}"))
        cut = str(tmp_path / "cut.bam")
        open(cut, "wb").write(buf.tobytes()[: buf.size // 2])
        with pytest.raises(IRFError):          # truncated file: refused as by a single-device context, and the context stays usable
            c.run_bam(cut)
        bad = bytearray(buf.tobytes())
        offs = _bgzf_block_offsets(bad)
        nxt = offs[len(offs) * 9 // 10 + 1]
        bad[nxt - 8] ^= 0x40                   # wrong CRC-32 footer in a block of the last shard: reported with the shard it was found in
        with pytest.raises(IRFError) as ei:
            c.run_bytes(bytes(bad))
        assert "CRC" in str(ei.value) and (len(devs) == 1 or "shard" in str(ei.value)), str(ei.value)
        c.run_bam(tiny["sorted"])
        c.write_outputs(str(tmp_path / f"again{k}"))
        assert_same_outputs(tiny["oracle_out"], str(tmp_path / f"again{k}"))
        c.close()


def test_multi_device_context_stranded_and_ragged_inputs(built, tiny, tmp_path):
    """The IR-dir table of a stranded library is merged like the nondir one; a header that spans several BGZF blocks is never cut;
    a BAM without alignments and a BAM with fewer blocks than devices run on fewer devices."""
    from bamutil import record, write_bam
    from irfinder_b200 import IRFinderCore
    meta = _gen(tmp_path, "s", genome="tiny", introns=300, pairs=40000, readlen=100, seed=12, stranded=-1, chunks=5)
    ostd = run_oracle(built, meta, str(tmp_path / "o"))
    c = IRFinderCore([0, 0, 0], stream_chunk_kb=64)
    c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
    c.run_bam(meta["sorted"])
    c.write_outputs(str(tmp_path / "g"))
    assert c.directionality == -1
    assert_same_outputs(str(tmp_path / "o"), str(tmp_path / "g"))
    assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(ostd)
    c.close()

    refs = [("chrA", 400000), ("chrB", 300000), ("chrEmpty", 50000)]
    many = refs + [("scaffold_%04d_random_with_a_long_name" % i, 1000 + i) for i in range(4000)]
    junc = [l.split("\t") for l in open(tiny["sj"]).read().split("\n") if l][:200]
    recs = []
    for i, j in enumerate(junc):
        ref, s, e = (0 if j[0] == "chrA" else 1), int(j[1]), int(j[2])
        recs.append(record("m%04d" % i, ref, s - 40, f"40M{e - s}N40M", 16 * (i & 1)))
        recs.append(record("s%04d" % i, 3 + (i * 19) % 4000, 10, "50M", 0))
    big = str(tmp_path / "many.bam")
    raw = write_bam(big, many, recs, level=1, block=3000)
    assert raw.index(b"m0000\0") > 2 * 3000
    empty = str(tmp_path / "empty.bam")
    write_bam(empty, refs, [], level=1)
    for name, bam in (("many", big), ("empty", empty)):
        m2 = dict(tiny, sorted=bam, collated=bam)
        ostd = run_oracle(built, m2, str(tmp_path / f"o_{name}"))
        c = IRFinderCore([0] * 4)
        c.load_ref(tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"])
        c.run_bam(bam)
        c.write_outputs(str(tmp_path / f"g_{name}"))
        assert_same_outputs(str(tmp_path / f"o_{name}"), str(tmp_path / f"g_{name}"))
        assert _stats_from_stdout(c.stdout_text) == _stats_from_stdout(ostd)
        c.close()


def test_cli_with_a_device_list(built, tiny, tmp_path):
    """IRF_DEVICES makes the reference invocation use several devices (here: three shards on GPU 0, and every GPU of the box)."""
    import torch
    exe = os.path.join(ROOT, "bin", "irfinder")
    lists = ["0,0,0", "0-%d" % (torch.cuda.device_count() - 1)]
    for k, devs in enumerate(lists):
        out = str(tmp_path / f"cli{k}")
        os.makedirs(out)
        r = subprocess.run([exe, out, tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"], tiny["sorted"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                           env=dict(os.environ, IRF_DEVICES=devs), timeout=300)
        assert r.returncode == 0, r.stderr
        assert_same_outputs(tiny["oracle_out"], out)
        assert _stats_from_stdout(r.stdout) == _stats_from_stdout(tiny["oracle_stdout"])
    bad = subprocess.run([exe, str(tmp_path), tiny["cover"], tiny["sj"], tiny["rc"], tiny["roi"], tiny["sorted"]], stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=dict(os.environ, IRF_DEVICES="0;1"))
    assert bad.returncode != 0


# ---- adjacent-mate path (name-collated input / STAR's pipe pairs without the hash table) ----

def test_adjacent_mates_pair_without_the_hash_table(built, tiny, tmp_path):
    """Name-collated input (header not SO:coordinate) takes the adjacent-mate path by default: nearly every pair is formed from two
    neighbouring records (only mates split by a chunk boundary go through the table); switching the path off, forcing it on for the
    coordinate-sorted copy (where only overlapping mates happen to be neighbours) and cutting the file into many small chunks all
    give the oracle's bytes - the path is a matter of speed, never of result."""
    want = _stats_from_stdout(tiny["oracle_stdout"])
    pairs = int(want["Total pairs processed"])
    runs = [("collated", {}, lambda a: a > 0.95 * pairs),
            ("collated", dict(adjacent_pairs=0), lambda a: a == 0),
            ("collated", dict(stream_chunk_kb=64), lambda a: 0.8 * pairs < a <= pairs),
            ("collated", dict(stream_pipeline=0), lambda a: a > 0.95 * pairs),
            ("sorted", {}, lambda a: a == 0),
            ("sorted", dict(adjacent_pairs=1), lambda a: 0 < a < pairs),
            ("sorted", dict(adjacent_pairs=1, stream_chunk_kb=64), lambda a: 0 < a < pairs)]
    for i, (which, opts, ok) in enumerate(runs):
        c = _gpu_run(tiny, str(tmp_path / f"g{i}"), bam=tiny[which], **opts)
        assert_same_outputs(tiny["oracle_out"], str(tmp_path / f"g{i}"))
        assert _stats_from_stdout(c.stdout_text) == want
        adj = c.timing()["adjacent_pairs"]
        assert ok(adj), (which, opts, adj, pairs)


def test_adjacent_path_on_runs_of_identical_mates(built, tiny, tmp_path):
    """Runs of 2, 3, 4 and 5 records that are all mates of each other (same name, mirrored positions - duplicates as some pipelines emit
    them): the first two pair as neighbours, the rest through the table, an odd one is left over as an error read, exactly as the
    reference's one-at-a-time loop does with identical records.  Also a pair whose halves are separated by one unrelated record."""
    from bamutil import record, write_bam
    refs = [("chrA", 400000), ("chrB", 300000), ("chrEmpty", 50000)]
    junc = [l.split("\t") for l in open(tiny["sj"]).read().split("\n") if l]
    junc = [(0 if j[0] == "chrA" else 1, int(j[1]), int(j[2])) for j in junc if j[0] in ("chrA", "chrB")][:40]
    P, PROPER, REV, MREV, R1, R2 = 0x1, 0x2, 0x10, 0x20, 0x40, 0x80
    recs = []
    for i, (ref, s, e) in enumerate(junc):
        gap = e - s
        a = record("run%03d" % i, ref, s - 30, f"30M{gap}N40M", P | PROPER | R1 | MREV, ref, s - 30, 0)      # level mates: pos == next_pos, so every copy is a mate of every other
        n = 2 + i % 4
        recs += [a] * n
        l = record("gap%03d" % i, ref, s - 60, "50M", P | PROPER | R1 | MREV, ref, e + 5, 300)
        r = record("gap%03d" % i, ref, e + 5, "50M", P | PROPER | R2 | REV, ref, s - 60, -300)
        recs += [l, record("mid%03d" % i, ref, s - 10, "10M", 0), r]
    bam = str(tmp_path / "runs.bam")
    write_bam(bam, refs, recs, level=1, block=2000)
    meta = dict(tiny, sorted=bam, collated=bam)
    ostd = run_oracle(built, meta, str(tmp_path / "o"))
    want = _stats_from_stdout(ostd)
    assert int(want["Error reads"]) == sum(1 for i in range(len(junc)) if (2 + i % 4) % 2), want
    for i, opts in enumerate([{}, dict(adjacent_pairs=0), dict(stream_chunk_kb=64)]):
        c = _gpu_run(meta, str(tmp_path / f"g{i}"), **opts)
        assert_same_outputs(str(tmp_path / "o"), str(tmp_path / f"g{i}"))
        assert _stats_from_stdout(c.stdout_text) == want
        if opts.get("adjacent_pairs", -1) != 0:
            assert c.timing()["adjacent_pairs"] >= len(junc)
