#!/usr/bin/env python
"""Full-size check of the segmented stream input (test infrastructure, run on the GPU box):

    python tests/stream_full.py [--config C2] [--segment-mb 1024] --out gpurun_out/stream_full.json

generates the coordinate-sorted BAM of a BASELINE config, counts it once with irf_run_bam (the file path) and once with
irf_run_stream from a pipe (`cat file |`, BGZF bytes, bounded segments, waiting mates carried between segments), and compares
the six output files byte for byte.  The irf_run_bam outputs of configs[1..3] are themselves compared with the CPU restatement
by tests/full_parity.py."""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
FILES = ["IRFinder-IR-nondir.txt", "IRFinder-IR-dir.txt", "IRFinder-JuncCount.txt", "IRFinder-SpansPoint.txt", "IRFinder-ChrCoverage.txt", "IRFinder-ROI.txt"]


def shas(d):
    return {n: hashlib.sha256(open(os.path.join(d, n), "rb").read()).hexdigest() for n in FILES if os.path.exists(os.path.join(d, n))}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2"); ap.add_argument("--segment-mb", type=int, default=1024); ap.add_argument("--workdir", default="/tmp/irf_stream_full")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "stream_full.json")); ap.add_argument("--scale", type=float, default=1.0)
    a = ap.parse_args()
    from irfinder_b200 import IRFinderCore, synth
    kw = dict(synth.CONFIGS[a.config]); kw["pairs"] = max(1000, int(kw["pairs"] * a.scale))
    meta = synth.generate(a.workdir, collated=False, chunks=256, **kw)
    core = IRFinderCore(0, stream_segment_mb=a.segment_mb)
    core.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
    t = time.perf_counter(); core.run_bam(meta["sorted"]); t_file = time.perf_counter() - t
    core.write_outputs(os.path.join(a.workdir, "out_file")); st_file = core.stats()
    res = {"config": a.config, "generator": kw, "bam_bytes": os.path.getsize(meta["sorted"]), "segment_mb": a.segment_mb, "run_bam_s": round(t_file, 3)}
    for rep in range(2):
        r, w = os.pipe()
        feeder = subprocess.Popen(["cat", meta["sorted"]], stdout=w); os.close(w)
        t = time.perf_counter(); core.run_stream(r, 0); dt = time.perf_counter() - t
        os.close(r); feeder.wait()
        res["run_stream_s_rep%d" % rep] = round(dt, 3)
    core.write_outputs(os.path.join(a.workdir, "out_stream"))
    sa, sb = shas(os.path.join(a.workdir, "out_file")), shas(os.path.join(a.workdir, "out_stream"))
    res.update(identical=sa == sb and st_file == core.stats(), sha256_file=sa, sha256_stream=sb, stats=core.stats(),
               segments=-(-res["bam_bytes"] // (a.segment_mb << 20)), stream_GBps=round(res["bam_bytes"] / res["run_stream_s_rep1"] / 1e9, 2))
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    json.dump(res, open(a.out, "w"), indent=1)
    print(json.dumps({k: res[k] for k in ("identical", "run_bam_s", "run_stream_s_rep0", "run_stream_s_rep1", "segments", "stream_GBps")}))
    sys.exit(0 if res["identical"] else 1)
