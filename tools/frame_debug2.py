"""Developer probe: framing repair rounds on the first generator chunks of a configuration (a BAM without its EOF block)."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from irfinder_b200 import IRFinderCore, synth
kw = dict(synth.CONFIGS[sys.argv[1]])
meta = synth.generate("/tmp/fd2_" + sys.argv[1], collated=False, chunks=256, chunk_range=(0, int(sys.argv[2])), **kw)
n = os.path.getsize(meta["sorted"]); host = torch.empty(n + 64, dtype=torch.uint8, pin_memory=True); hv = host.numpy()
with open(meta["sorted"], "rb") as f: f.readinto(memoryview(hv)[:n])
c = IRFinderCore(0, require_eof=0); c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
c.run_buffer(hv.ctypes.data, n)
print(c.stats(), c.timing())
