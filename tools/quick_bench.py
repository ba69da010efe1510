"""Developer timing loop (not the contract bench): stage once, then time count+finalize for several inflate variants."""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from irfinder_b200 import IRFinderCore, synth

ap = argparse.ArgumentParser()
ap.add_argument("--pairs", type=int, default=1_000_000); ap.add_argument("--genome", default="chr22"); ap.add_argument("--introns", type=int, default=5000)
ap.add_argument("--readlen", type=int, default=100); ap.add_argument("--variants", default="0,1,2,3,4,5,6"); ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--ctas", default="0"); ap.add_argument("--de-percent", type=int, default=0)
ap.add_argument("--out", default="/tmp/qb"); ap.add_argument("--single", action="store_true")
ap.add_argument("--depth-tiles", default="0"); ap.add_argument("--intronic", type=float, default=0.05); ap.add_argument("--hot-introns", type=int, default=0)
a = ap.parse_args()
t0 = time.time()
meta = synth.generate(a.out, genome=a.genome, introns=a.introns, pairs=a.pairs, readlen=a.readlen, seed=2, collated=False, single=a.single, intronic=a.intronic, hot_introns=a.hot_introns)
print("generated in %.1fs" % (time.time() - t0), {k: meta[k] for k in ("records", "inflated_bytes", "compressed_bytes")}, flush=True)
data = np.fromfile(meta["sorted"], dtype=np.uint8)
for v, ctas, dtile in [(int(x), int(y), int(z)) for x in a.variants.split(",") for y in a.ctas.split(",") for z in a.depth_tiles.split(",")]:
    c = IRFinderCore(0, inflate_variant=v, inflate_ctas=ctas, de_percent=a.de_percent, resident_pipeline=0, **({"depth_tile": dtile} if dtile else {}))
    c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
    c.stage(data.ctypes.data, data.size)
    best = None
    for r in range(a.reps):
        c.reset(); t = time.time(); c.count(); c.finalize(); wall = time.time() - t
        tm = c.timing(); tm["wall_ms"] = wall * 1e3
        if best is None or tm["inflate_kernel_ms"] < best["inflate_kernel_ms"]: best = tm
    gbs = (meta["inflated_bytes"] + meta["compressed_bytes"]) / best["inflate_kernel_ms"] / 1e6
    print(json.dumps(dict(variant=v, ctas=ctas, depth_tile=dtile, inflate_GBps=round(gbs, 1), **{k: round(x, 3) for k, x in best.items()})), flush=True)
    c.close()
