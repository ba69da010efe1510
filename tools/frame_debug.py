"""Developer probe: repair rounds of the record framing on a generated workload (IRF_DEBUG_FRAMING=1 prints them)."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from irfinder_b200 import IRFinderCore, synth
kw = dict(synth.CONFIGS[sys.argv[1]]); kw["pairs"] = int(sys.argv[2])
meta = synth.generate("/tmp/fd_" + sys.argv[1], collated=False, **kw)
n = os.path.getsize(meta["sorted"]); host = torch.empty(n + 64, dtype=torch.uint8, pin_memory=True); hv = host.numpy()
with open(meta["sorted"], "rb") as f: f.readinto(memoryview(hv)[:n])
c = IRFinderCore(0, **({"stream_chunk_kb": int(sys.argv[3])} if len(sys.argv) > 3 else {})); c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
c.run_buffer(hv.ctypes.data, n)
print(c.stats(), c.timing())
