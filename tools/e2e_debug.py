import sys, os, numpy as np, torch
sys.path.insert(0, "/root/repo")
from irfinder_b200 import IRFinderCore, synth
meta = synth.generate("/tmp/qb", genome="grch38", introns=200000, pairs=int(os.environ.get("PAIRS", "20000000")), readlen=150, seed=2, collated=False)
data = np.fromfile(meta["sorted"], dtype=np.uint8)
host = torch.empty(data.size + 64, dtype=torch.uint8, pin_memory=True); hv = host.numpy(); hv[:data.size] = data
c = IRFinderCore(0); c.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
for i in range(3):
    print("---- run", i, file=sys.stderr, flush=True)
    c.run_buffer(hv.ctypes.data, data.size); c.write_outputs("/dev/shm/e2e_dbg")
print(c.timing())
print({f: os.path.getsize("/dev/shm/e2e_dbg/" + f) for f in os.listdir("/dev/shm/e2e_dbg")})
