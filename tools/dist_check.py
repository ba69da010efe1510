"""Multi-GPU correctness check (run under torchrun): sharded NCCL run vs the CPU oracle, byte for byte.
   python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tools/dist_check.py [--pairs P]"""
import argparse, os, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from irfinder_b200 import IRFinderCore, build, synth
from irfinder_b200.distributed import run_sharded

ap = argparse.ArgumentParser(); ap.add_argument("--pairs", type=int, default=200000); ap.add_argument("--genome", default="chr22"); ap.add_argument("--introns", type=int, default=3000)
a = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dist.init_process_group("nccl", device_id=torch.device("cuda", local))
wd = "/tmp/irf_dist_check"
if rank == 0:
    meta = synth.generate(wd, genome=a.genome, introns=a.introns, pairs=a.pairs, readlen=100, seed=77, chunks=16)
dist.barrier()
meta = synth.describe(wd)
core = IRFinderCore(local); core.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
buf = np.fromfile(meta["sorted"], dtype=np.uint8)
info = run_sharded(core, buf, rank, world, device=torch.device("cuda", local))
out = f"{wd}/gpu{rank}"; core.write_outputs(out)
ok = True
if rank == 0:
    oracle = build.build_oracle(); os.makedirs(f"{wd}/cpu", exist_ok=True)
    subprocess.run([oracle, f"{wd}/cpu", meta["cover"], meta["sj"], meta["rc"], meta["roi"], meta["collated"]], check=True, stdout=subprocess.DEVNULL)
dist.barrier()
for name in sorted(os.listdir(f"{wd}/cpu")):
    same = open(f"{wd}/cpu/{name}", "rb").read() == open(f"{out}/{name}", "rb").read()
    ok &= same
    if not same: print(f"rank {rank}: {name} DIFFERS")
print(f"rank {rank}/{world}: {'IDENTICAL to oracle' if ok else 'MISMATCH'}; halo records {info.get('halo_records')}, recount {info['recount']}, stats {core.stats()}", flush=True)
dist.barrier(); dist.destroy_process_group()
sys.exit(0 if ok else 1)
