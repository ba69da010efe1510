"""In-process multi-device context at a real size: one BAM, counted by irf_run_bam on 1 device and on every device list given,
wall-clock per run and sha256 of the six output files (they must not depend on the device list).

    python tools/multi_device_check.py --pairs 20000000 --lists 0 0,1 0,0 --out gpurun_out/multi_device.json
"""
import argparse, hashlib, json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from irfinder_b200 import IRFinderCore, synth

ap = argparse.ArgumentParser()
ap.add_argument("--pairs", type=int, default=20_000_000); ap.add_argument("--genome", default="grch38"); ap.add_argument("--introns", type=int, default=200_000)
ap.add_argument("--readlen", type=int, default=150); ap.add_argument("--lists", nargs="+", default=["0", "0,0"]); ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--work", default="/tmp/mdc"); ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "multi_device.json")); ap.add_argument("--cli", type=int, default=1)
ap.add_argument("--buffer", type=int, default=1, help="also time irf_run_buffer on the file in page-locked host memory")
a = ap.parse_args()


def sha_dir(d):
    h = hashlib.sha256()
    for name in sorted(os.listdir(d)):
        h.update(name.encode()); h.update(open(os.path.join(d, name), "rb").read())
    return h.hexdigest()


t0 = time.time()
meta = synth.generate(a.work, genome=a.genome, introns=a.introns, pairs=a.pairs, readlen=a.readlen, seed=2, collated=False, chunks=64)
res = dict(pairs=a.pairs, compressed_bytes=meta["compressed_bytes"], generation_s=round(time.time() - t0, 1), runs=[])
subprocess.run(["cat", meta["sorted"]], stdout=subprocess.DEVNULL)      # page cache
pinned = None
if a.buffer:
    import numpy as np, torch
    n = os.path.getsize(meta["sorted"])
    pinned = torch.empty(n, dtype=torch.uint8, pin_memory=True)
    pinned.numpy()[:] = np.fromfile(meta["sorted"], dtype=np.uint8)
for l in a.lists:
    devs = [int(x) for x in l.split(",")]
    c = IRFinderCore(devs)
    c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
    walls = []
    for r in range(a.reps):
        t = time.time(); c.run_bam(meta["sorted"]); walls.append(round((time.time() - t) * 1e3, 1))
    out = os.path.join(a.work, "out_" + l.replace(",", "_")); c.write_outputs(out)
    bwalls = []
    if pinned is not None:
        for r in range(a.reps):
            t = time.time(); c.run_buffer(pinned.data_ptr(), pinned.numel()); bwalls.append(round((time.time() - t) * 1e3, 1))
        outb = out + "_buf"; c.write_outputs(outb)
        assert sha_dir(outb) == sha_dir(out), "irf_run_buffer and irf_run_bam disagree"
    rec = dict(devices=devs, run_bam_ms=walls, run_buffer_pinned_ms=bwalls, best_Mpairs_per_s=round(a.pairs / min(walls) / 1e3, 1), sha256=sha_dir(out), stats=c.stats())
    c.close()
    if a.cli:
        out2 = out + "_cli"; os.makedirs(out2, exist_ok=True)
        t = time.time()
        p = subprocess.run([os.path.join(ROOT, "bin", "irfinder"), out2, meta["cover"], meta["sj"], meta["rc"], meta["roi"], meta["sorted"]], env=dict(os.environ, IRF_DEVICES=l), stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        rec["cli_wall_s"] = round(time.time() - t, 2); rec["cli_rc"] = p.returncode; rec["cli_sha256"] = sha_dir(out2) if p.returncode == 0 else p.stderr[-300:]
    res["runs"].append(rec)
    print(json.dumps(rec), flush=True)
res["identical"] = len({r["sha256"] for r in res["runs"]} | {r.get("cli_sha256", res["runs"][0]["sha256"]) for r in res["runs"]}) == 1
os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
json.dump(res, open(a.out, "w"), indent=1)
print("identical:", res["identical"])
