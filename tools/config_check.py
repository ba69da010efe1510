"""Runs BASELINE.json configs[2] / configs[3]-shaped workloads end to end once and prints counters, timings and basic consistency checks."""
import sys, os, time, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from irfinder_b200 import IRFinderCore, synth
which = sys.argv[1]; scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
kw = dict(synth.CONFIGS[which]); kw["pairs"] = int(kw["pairs"] * scale)
t = time.time(); meta = synth.generate("/tmp/cfg_" + which, collated=False, **kw); print(which, "generated", kw["pairs"], "in %.1fs" % (time.time() - t), flush=True)
n = os.path.getsize(meta["sorted"]); host = torch.empty(n + 64, dtype=torch.uint8, pin_memory=True); hv = host.numpy()
with open(meta["sorted"], "rb") as f: f.readinto(memoryview(hv)[:n])
c = IRFinderCore(0); c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
for i in range(2):
    t = time.time(); c.run_buffer(hv.ctypes.data, n); c.write_outputs("/dev/shm/cfg_out_" + which); dt = time.time() - t
st = c.stats(); units = meta["pairs"] + meta["singles"]
print("e2e %.1f ms -> %.1f M fragments/s" % (dt * 1e3, units / dt / 1e6), "directionality", c.directionality, st)
print("checks:", st["pairs"] + st["singles"] == units, int(c.table("chr").astype(np.int64).sum()) == units, st["error"] == 0, st["records"] == meta["records"], "max depth", int(c.table("depth").max()))
print(sorted(os.listdir("/dev/shm/cfg_out_" + which)), c.timing())
