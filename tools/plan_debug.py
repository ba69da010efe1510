"""Debug: run_plan with world=1 (no torch.distributed) at a given size, lazy staging."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from irfinder_b200 import IRFinderCore, synth
from irfinder_b200.distributed import ShardPlan, run_plan
pairs = int(sys.argv[1]); mode = sys.argv[2] if len(sys.argv) > 2 else "plan"
meta = synth.generate("/tmp/pd", genome="grch38", introns=200000, pairs=pairs, readlen=150, seed=2, collated=False)
n = os.path.getsize(meta["sorted"])
host = torch.empty(n + 64, dtype=torch.uint8, pin_memory=True); hv = host.numpy()
with open(meta["sorted"], "rb") as f: f.readinto(memoryview(hv)[:n])
c = IRFinderCore(0); c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
if mode == "plan":
    plan = ShardPlan(hv[: 1 << 26], hv[:n], n, 0, 1)
    for i in range(2):
        run_plan(c, plan, device=torch.device("cuda", 0)); print("plan run", i, c.timing()["inflate_wall_ms"], c.stats()["pairs"], flush=True)
else:
    c.set_option("stage_lazy", 1)
    for i in range(2):
        c.reset(); c.stage(hv.ctypes.data, n); c.count(); c.finalize(); print("lazy run", i, c.timing()["inflate_wall_ms"], c.stats()["pairs"], flush=True)
