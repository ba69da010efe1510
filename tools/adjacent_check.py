"""Times the adjacent-mate path (option adjacent_pairs) on a name-collated BAM: count + finalize on resident bytes, stage by stage
and through the resident pipeline, with the path off and on.

    python tools/adjacent_check.py --pairs 20000000 --out gpurun_out/adjacent_check.json
"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from irfinder_b200 import IRFinderCore, synth

ap = argparse.ArgumentParser()
ap.add_argument("--pairs", type=int, default=20_000_000); ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--work", default="/tmp/adj"); ap.add_argument("--out", default="gpurun_out/adjacent_check.json")
a = ap.parse_args()
meta = synth.generate(a.work, genome="grch38", introns=200_000, pairs=a.pairs, readlen=150, seed=2, sorted_bam=False, collated=True, chunks=64)
data = np.fromfile(meta["collated"], dtype=np.uint8)
res = dict(pairs=a.pairs, compressed_bytes=int(data.size), runs=[])
for pipeline in (0, 1):
    for adj in (0, -1):
        c = IRFinderCore(0, adjacent_pairs=adj, resident_pipeline=pipeline)
        c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
        c.stage(data.ctypes.data, data.size)
        best = None
        for r in range(a.reps):
            c.reset(); t = time.time(); c.count(); c.finalize(); wall = (time.time() - t) * 1e3
            tm = c.timing(); tm["wall_ms"] = wall
            if best is None or wall < best["wall_ms"]:
                best = tm
        rec = dict(resident_pipeline=pipeline, adjacent_pairs_option=adj, wall_ms=round(best["wall_ms"], 1), counting_ms=round(best["counting_ms"], 2), count_total_ms=round(best["count_total_ms"], 1),
                   pairs_by_adjacency=int(best["adjacent_pairs"]), pairs=c.stats()["pairs"])
        res["runs"].append(rec); print(json.dumps(rec), flush=True)
        c.close()
os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
json.dump(res, open(a.out, "w"), indent=1)
