"""Developer sweep: resident re-count (what bench.py times as `value`) for several engine splits of the overlapped pipeline."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from irfinder_b200 import IRFinderCore, synth

ap = argparse.ArgumentParser()
ap.add_argument("--pairs", type=int, default=20_000_000); ap.add_argument("--genome", default="grch38"); ap.add_argument("--introns", type=int, default=200000)
ap.add_argument("--readlen", type=int, default=150); ap.add_argument("--shares", default="-1,50,60,65,70,80,100"); ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--out", default="/tmp/ov"); ap.add_argument("--chunk-kb", type=int, default=0)
a = ap.parse_args()
t0 = time.time()
meta = synth.generate(a.out, genome=a.genome, introns=a.introns, pairs=a.pairs, readlen=a.readlen, seed=2, collated=False)
print("generated in %.1fs" % (time.time() - t0), {k: meta[k] for k in ("records", "inflated_bytes", "compressed_bytes")}, flush=True)
data = np.fromfile(meta["sorted"], dtype=np.uint8)
ref = None
for share in [int(x) for x in a.shares.split(",")]:
    opts = {"resident_pipeline": 0} if share < 0 else {"de_percent_overlap": share}
    if a.chunk_kb: opts["stream_chunk_kb"] = a.chunk_kb
    c = IRFinderCore(0, **opts)
    c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
    c.stage(data.ctypes.data, data.size)
    best = None
    for r in range(a.reps):
        c.reset(); t = time.time(); c.count(); c.finalize(); wall = (time.time() - t) * 1e3
        if best is None or wall < best[0]: best = (wall, c.timing())
    st = c.stats()
    if ref is None: ref = st
    print(json.dumps(dict(share=share, wall_ms=round(best[0], 1), same=(st == ref), Mpairs_s=round(a.pairs / best[0] / 1e3, 1), **{k: round(x, 1) for k, x in best[1].items()})), flush=True)
    c.close()
