"""Synthetic workloads of BASELINE.json's configs, produced by tools/irf_synth."""
import json
import os
import subprocess

from .build import ROOT, build_tools

# name -> generator options; seeds 1-5 as in SURVEY.md §8d
CONFIGS = {
    "C1": dict(genome="chr22", introns=5000, pairs=1_000_000, readlen=100, seed=1),
    "C2": dict(genome="grch38", introns=200_000, pairs=50_000_000, readlen=150, seed=2),
    "C3": dict(genome="grch38", introns=200_000, pairs=50_000_000, readlen=150, seed=3, stranded=-1),
    "C4": dict(genome="grch38", introns=200_000, pairs=100_000_000, readlen=75, seed=4, single=True, intronic=0.4, hot_introns=5000),
    "C5": dict(genome="grch38", introns=200_000, pairs=500_000_000, readlen=150, seed=5),
}


def generate(outdir, genome="chr22", introns=5000, pairs=100_000, readlen=100, seed=1, single=False, stranded=0, intronic=0.05,
             hot_introns=0, level=1, threads=0, chunks=256, sorted_bam=True, collated=True, align_records=False, refs_only=False, chunk_range=None):
    exe = build_tools()
    cmd = [exe, "--out", outdir, "--genome", genome, "--introns", str(introns), "--pairs", str(pairs), "--readlen", str(readlen), "--seed", str(seed),
           "--stranded", str(stranded), "--intronic", str(intronic), "--hot-introns", str(hot_introns), "--level", str(level), "--threads", str(threads), "--chunks", str(chunks)]
    if single:
        cmd.append("--single")
    if not sorted_bam:
        cmd.append("--no-sorted")
    if not collated:
        cmd.append("--no-collated")
    if align_records:
        cmd.append("--align-records")
    if refs_only:
        cmd.append("--refs-only")
    if chunk_range is not None:   # only the byte range of chunks [a,b) of the whole file (header with a == 0, EOF marker with b == chunks)
        cmd += ["--chunk-range", "%d:%d" % tuple(chunk_range)]
    os.makedirs(outdir, exist_ok=True)
    subprocess.run(cmd, check=True)
    return describe(outdir)


def describe(outdir):
    """Paths and counts of a generated workload directory."""
    meta = {}
    mp = os.path.join(outdir, "meta.json")
    if os.path.exists(mp):
        with open(mp) as f:
            meta = json.load(f)
    d = os.path.join(outdir, "ref", "IRFinder")
    meta.update(dir=outdir, ref_dir=os.path.join(outdir, "ref"), cover=os.path.join(d, "ref-cover.bed"), sj=os.path.join(d, "ref-sj.ref"),
                rc=os.path.join(d, "ref-read-continues.ref"), roi=os.path.join(d, "ref-ROI.bed"), sorted=os.path.join(outdir, "sorted.bam"),
                collated=os.path.join(outdir, "collated.bam"))
    return meta
