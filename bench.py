#!/usr/bin/env python
"""bench.py — BAM -> IR table throughput of the B200 path (BASELINE.json metric) and the CPU reference arm.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--pairs P]

A "step" is one complete pass of the hot path over the workload: BGZF inflate, record framing, mate
pairing, every ProcessBlocks, depth scan and per-intron statistics.
  value : million read pairs / s with the compressed BAM already resident in HBM (irf_reset + irf_count + irf_finalize)
  e2e   : the same through the reference-facing call with HOST buffers: irf_run_buffer (pinned host BGZF bytes ->
          device in the timed region) + irf_write_outputs (result tables device -> host, six text tables formatted)
  roofline : dominant kernel = inflate_tok_kernel; algorithmic bytes = compressed in + inflated out
  cpu_baseline : the CPU restatement in oracle/ (the mounted reference has no source) on one host core,
          bounded sample of the same workload
Workload: BASELINE.json configs[1] (50M pairs 2x150, GRCh38-shaped, ~200k introns); inputs (10 GB) exceed L2.
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "M read pairs/sec BAM->IR table"
UNIT = "M read pairs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--pairs", type=int, default=50_000_000, help="read pairs per GPU (configs[1] = 50M)")
    ap.add_argument("--introns", type=int, default=200_000)
    ap.add_argument("--genome", default="grch38")
    ap.add_argument("--readlen", type=int, default=150)
    ap.add_argument("--workdir", default=os.environ.get("IRF_BENCH_DIR", "/tmp/irf_bench"))
    ap.add_argument("--cpu-sample-pairs", type=int, default=1_000_000)
    ap.add_argument("--inflate-variant", type=int, default=-1)
    ap.add_argument("--de-percent-overlap", type=int, default=0, help="share of the file's chunks given to the decompression engine by the overlapped resident pipeline (library default when 0)")
    ap.add_argument("--resident-pipeline", type=int, default=-1, help="0 = count resident bytes stage by stage, 1 = overlapped pipeline when the input is large enough (library default), 2 = always")
    ap.add_argument("--numa-bind", type=int, default=0, help="1 = run this rank on the CPUs of its GPU's NUMA node (so its page-locked buffer is node-local); off by default, not yet measured")
    ap.add_argument("--skip-parity", type=int, default=0, help="1 = skip the configs[0] parity check through the NCCL path that a multi-GPU run does before its timed loops")
    ap.add_argument("--de-percent", type=int, default=-1, help="share of BGZF blocks given to the hardware decompression engine (library default when < 0)")
    return ap.parse_args()


def workload_name(a, total_pairs):
    return f"configs[1]-shaped: {total_pairs / 1e6:g}M pairs 2x{a.readlen} unstranded, coordinate-sorted BGZF BAM (zlib level 1), synthetic {a.genome} genome, {a.introns} target introns"


def gen_cached(outdir, **kw):
    """Generates a workload once per (directory, parameters); both arms and repeated runs on one box share it."""
    from irfinder_b200 import synth
    key = json.dumps(kw, sort_keys=True)
    kp = os.path.join(outdir, "key.json")
    if os.path.exists(kp) and open(kp).read() == key and os.path.exists(os.path.join(outdir, "meta.json")):
        return synth.describe(outdir)
    shutil.rmtree(outdir, ignore_errors=True)
    meta = synth.generate(outdir, **kw)
    with open(kp, "w") as f:
        f.write(key)
    return meta


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region (B200_PROFILING.md)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def oracle_run(oracle, meta, bam, outdir):
    os.makedirs(outdir, exist_ok=True)
    t = time.perf_counter()
    r = subprocess.run([oracle, outdir, meta["cover"], meta["sj"], meta["rc"], meta["roi"], bam], stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True)
    dt = time.perf_counter() - t
    if r.returncode != 0:
        raise RuntimeError("oracle failed: " + r.stderr[-500:])
    return dt


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(a):
    """CPU arm: the reference's own implementation is absent from /root/reference (README.md only), so this times the
    in-repo CPU restatement (oracle/irfinder_cpu, kind "port"), one single-threaded process per host core, each on its
    own region-shaped shard of the same workload, gzip inflate included (as the reference's `gzip -cd |` pipe is)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from irfinder_b200 import build
    oracle = build.build_oracle()
    build.build_tools()
    cores = os.cpu_count() or 1
    shard_pairs = max(100_000, min(2_000_000, a.pairs // max(cores, 1)))
    wd = os.path.join(a.workdir, "ref_arm")
    metas = []
    t0 = time.perf_counter()
    for s in range(cores):  # one independent BAM (own seed) per host core
        m = gen_cached(os.path.join(wd, f"s{s}"), genome=a.genome, introns=a.introns, pairs=shard_pairs, readlen=a.readlen, seed=2 + s, sorted_bam=False, chunks=16)
        metas.append(m)
    gen_s = time.perf_counter() - t0
    pairs_total = sum(m["pairs"] + m["singles"] for m in metas)

    def one_step():
        res = [None] * cores

        def work(i):
            res[i] = oracle_run(oracle, metas[i], metas[i]["collated"], os.path.join(wd, f"out{i}"))
        th = [threading.Thread(target=work, args=(i,)) for i in range(cores)]
        t = time.perf_counter()
        [x.start() for x in th]
        [x.join() for x in th]
        return time.perf_counter() - t
    for _ in range(a.warmup):
        one_step()
    times = [one_step() for _ in range(a.steps)]
    dt = sum(times)
    value = pairs_total * a.steps / dt / 1e6
    sample = f"{cores} processes x {shard_pairs / 1e6:g}M-pair shards of the workload per step (collated BAM, gzip inflate + reference load included)"
    line = {"metric": METRIC, "value": round(value, 4), "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(dt / a.steps * 1e3, 2),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32/u32 counts + f64 ratios", "data": "synthetic", "impl": "reference",
            "config": {"workload": workload_name(a, a.pairs * a.gpus), "sampled": sample, "generation_s": round(gen_s, 1)},
            "cpu_baseline": {"value": round(value, 4), "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": round(value, 4), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    emit(line)


# ------------------------------------------------------------------------------------------------ B200 arm
OUT_FILES = ["IRFinder-IR-nondir.txt", "IRFinder-IR-dir.txt", "IRFinder-JuncCount.txt", "IRFinder-SpansPoint.txt", "IRFinder-ChrCoverage.txt", "IRFinder-ROI.txt"]


def sha_dir(d):
    import hashlib
    out = {}
    for name in OUT_FILES:
        p = os.path.join(d, name)
        if os.path.exists(p):
            out[name] = hashlib.sha256(open(p, "rb").read()).hexdigest()
    return out


def nccl_parity_check(rank, world, local, dist):
    """Before the timed loops of a multi-GPU run: BASELINE configs[0] (1M pairs 2x100 over a 50 Mb chr22-like genome) goes through the
    real NCCL path (run_sharded over this process group) and rank 0 compares the six output files with the CPU restatement's
    (the oracle as the checker, on the name-collated copy of the same reads).  Returns ("identical" | first differing file, sha256s)."""
    import numpy as np
    import torch
    from irfinder_b200 import IRFinderCore, build, synth
    from irfinder_b200.distributed import run_sharded
    wd = os.path.join("/dev/shm" if os.path.isdir("/dev/shm") else "/tmp", "irf_bench_parity")
    if rank == 0:
        meta = gen_cached(wd, **synth.CONFIGS["C1"])
    dist.barrier()
    meta = synth.describe(wd)
    buf = np.fromfile(meta["sorted"], dtype=np.uint8)
    core = IRFinderCore(local)
    core.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
    info = run_sharded(core, buf, rank, world, device=torch.device("cuda", local))
    verdict, shas = None, {}
    if rank == 0:
        got = os.path.join(wd, "gpu_out")
        shutil.rmtree(got, ignore_errors=True)
        core.write_outputs(got)
        want = os.path.join(wd, "oracle_out")
        if not os.path.exists(os.path.join(want, "done")):
            shutil.rmtree(want, ignore_errors=True)
            oracle_run(build.build_oracle(), meta, meta["collated"], want)
            open(os.path.join(want, "done"), "w").close()
        verdict = "identical"
        for name in OUT_FILES:
            pa, pb = os.path.join(want, name), os.path.join(got, name)
            if os.path.exists(pa) != os.path.exists(pb) or (os.path.exists(pa) and open(pa, "rb").read() != open(pb, "rb").read()):
                verdict = name
                break
        shas = sha_dir(got)
    core.close()
    dist.barrier()
    return verdict, shas, info.get("foreign_depth_entries")


def run_b200(a):
    import numpy as np
    import torch
    from irfinder_b200 import IRFinderCore, build
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if a.numa_bind:
        numa_bind(local)
    parity_nccl, parity_sha, parity_foreign = None, {}, None
    if world > 1 and not a.skip_parity:
        if rank == 0:
            build.build_tools()
        parity_nccl, parity_sha, parity_foreign = nccl_parity_check(rank, world, local, dist)
    total_pairs = a.pairs * world
    base = a.workdir
    if world > 1 and os.path.isdir("/dev/shm"):   # N shards are generated at once: keep them off a possibly small disk
        try:
            st = os.statvfs("/dev/shm")
            if st.f_bavail * st.f_frsize > 1.5 * world * a.pairs * 300:   # all ranks' parts exist at once (~290 B per pair) until each has been read into pinned memory
                base = "/dev/shm/irf_bench"
        except OSError:
            pass
    wd = os.path.join(base, f"w{world}_r{rank}")
    t0 = time.perf_counter()
    if rank == 0:
        build.build_tools()
    if world > 1:
        dist.barrier()
    # One logical BAM of `total_pairs`; with N ranks each rank generates ONLY its own byte range of that file
    # (`irf_synth --chunk-range`: the concatenation of the parts is byte-identical to the whole file), so no rank
    # ever holds more than its shard and generation parallelises across the host cores.
    nchunks = max(256, 64 * world)
    clo, chi = nchunks * rank // world, nchunks * (rank + 1) // world
    threads = max(1, (os.cpu_count() or 1) // world)
    meta = gen_cached(wd, genome=a.genome, introns=a.introns, pairs=total_pairs, readlen=a.readlen, seed=2, collated=False, chunks=nchunks,
                      chunk_range=(clo, chi) if world > 1 else None, threads=threads)
    gen_s = time.perf_counter() - t0
    fsize = os.path.getsize(meta["sorted"])
    host = torch.empty(fsize + (1 << 18) + 64, dtype=torch.uint8, pin_memory=True)   # pinned host copy of this rank's BGZF bytes (+ room for the look-ahead)
    hv = host.numpy()
    with open(meta["sorted"], "rb") as f:
        f.readinto(memoryview(hv)[:fsize])
    if world > 1:   # N shards of this size would otherwise sit on disk at once; the bytes now live in pinned memory
        hdr_bytes = None if rank == 0 else np.fromfile(meta["sorted"] + ".header", dtype=np.uint8)
        for fn in (meta["sorted"], os.path.join(wd, "key.json")):
            if os.path.exists(fn):
                os.remove(fn)
    counts = torch.tensor([meta["pairs"] + meta["singles"], meta["records"], meta["compressed_bytes"], meta["inflated_bytes"]], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(counts)
    n_units, n_records, comp_bytes, infl_bytes = (int(x) for x in counts.tolist())   # fragments = pairs (a few orphans count as one unit each)
    opts = {} if a.inflate_variant < 0 else {"inflate_variant": a.inflate_variant}
    if a.de_percent >= 0:
        opts["de_percent"] = a.de_percent
    if a.de_percent_overlap > 0:
        opts["de_percent_overlap"] = a.de_percent_overlap
    if a.resident_pipeline >= 0:
        opts["resident_pipeline"] = a.resident_pipeline
    core = IRFinderCore(local, **opts)
    t1 = time.perf_counter()
    core.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
    ref_s = time.perf_counter() - t1

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---------------- value: inputs resident in HBM
    inflate_ms, sm_inflate_ms, launches, inflate_launches, stage_t = [], [], 0, 0, {}
    if world == 1:
        core.stage(hv.ctypes.data, fsize)

        def step():
            core.reset(); core.count(); core.finalize()
    else:
        from irfinder_b200.distributed import ShardPlan, run_plan
        plan = ShardPlan.from_parts(hv[:fsize], hdr_bytes, rank, world, dist, torch.device("cuda", local))

        merge_phases = {}

        def step():   # the shard is copied to HBM by the first (warm-up) call and stays resident, as at N = 1
            info = run_plan(core, plan, device=torch.device("cuda", local), restage=False)
            merge_phases.clear(); merge_phases.update(info.get("phase_ms", {})); merge_phases["foreign_depth_entries"] = info.get("foreign_depth_entries")
    for _ in range(a.warmup):
        step()
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    t = time.perf_counter()
    for _ in range(a.steps):
        step()
        tm = core.timing()
        inflate_ms.append(tm["inflate_kernel_ms"]); sm_inflate_ms.append(tm.get("sm_inflate_ms", 0.0)); launches += int(tm["launches"]); inflate_launches += int(tm["inflate_launches"]); stage_t = tm
    barrier()
    dt = time.perf_counter() - t
    if world > 1:
        v = torch.tensor([dt], dtype=torch.float64, device="cuda"); dist.all_reduce(v, op=dist.ReduceOp.MAX); dt = float(v.item())
    value = n_units * a.steps / dt / 1e6
    stats = core.stats()
    # size-independent consistency checks at full size (the byte-level parity tests run at sizes the CPU oracle can handle)
    chk = core.table("chr").astype(np.int64).sum()
    checks = {"fragments == pairs + singles": int(stats["pairs"] + stats["singles"]) == n_units, "sum(ChrCoverage) == fragments": int(chk) == n_units,
              "long + short + intersect == pairs": stats["long"] + stats["short"] + stats["intersect"] == stats["pairs"], "no error reads": stats["error"] == 0,
              "records == generated": stats["records"] == n_records}

    # ---------------- e2e: host buffer in, six text tables out
    outdir = os.path.join("/dev/shm" if os.path.isdir("/dev/shm") else wd, f"irf_bench_out_{rank}")
    if world == 1:
        def e2e_step():
            core.run_buffer(hv.ctypes.data, fsize); core.write_outputs(outdir)
    else:
        def e2e_step():
            run_plan(core, plan, device=torch.device("cuda", local))
            if rank == 0:
                core.write_outputs(outdir)
    for _ in range(min(a.warmup, 2)):
        e2e_step()
    barrier()
    t = time.perf_counter()
    for _ in range(a.steps):
        e2e_step()
    barrier()
    dte = time.perf_counter() - t
    e2e_stage = core.timing()
    if world > 1:
        v = torch.tensor([dte], dtype=torch.float64, device="cuda"); dist.all_reduce(v, op=dist.ReduceOp.MAX); dte = float(v.item())
    e2e_value = n_units * a.steps / dte / 1e6
    # ---------------- e2e from the FILE (SURVEY.md §8d t_e2e: open(bam) -> six files closed), N = 1: irf_run_bam reads the BAM (page cache)
    # through its ring of page-locked buffers; reported beside `e2e`, whose input already sits in page-locked memory
    e2e_file = None
    if world == 1:
        nf = max(1, min(a.steps, 3))
        core.run_bam(meta["sorted"]); core.write_outputs(outdir)
        torch.cuda.synchronize()
        t = time.perf_counter()
        for _ in range(nf):
            core.run_bam(meta["sorted"]); core.write_outputs(outdir)
        dtf = time.perf_counter() - t
        e2e_file = {"value": round(n_units * nf / dtf / 1e6, 3), "unit": UNIT, "ms_per_step": round(dtf / nf * 1e3, 3), "steps": nf,
                    "what": "irf_run_bam(path) + irf_write_outputs: the BGZF file is read from the page cache through the library's ring of page-locked buffers inside the timed region"}
    clocks = sampler.stop() if sampler else None
    d2h = sum(core.table(x).nbytes for x in ("rows_nondir", "rows_dir", "spans", "roi", "chr", "junctions", "stats"))
    out_bytes = sum(os.path.getsize(os.path.join(outdir, f)) for f in os.listdir(outdir)) if (rank == 0 and os.path.isdir(outdir)) else 0

    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        return
    # ---------------- roofline of the dominant kernel (inflate): algorithmic bytes = compressed in + inflated out
    peak, peak_src = measured_peak()
    alg_bytes = (comp_bytes + infl_bytes) / world   # per GPU
    avg_inflate_ms = sum(inflate_ms) / max(len(inflate_ms), 1)
    stage_achieved = alg_bytes / (avg_inflate_ms * 1e-3) / 1e9 if avg_inflate_ms > 0 else 0.0
    # Inflate runs on two engines at once: one persistent inflate_kernel launch takes the leading part of the file, the
    # Blackwell hardware decompression engine the rest, and framing + counting follow chunk by chunk behind them.  The
    # roofline entry is the KERNEL's: `achieved` = the algorithmic bytes (compressed in + inflated out) of the blocks it
    # inflated / its own launch duration, both read from the library's timing table (CUDA events on its stream).  The figure
    # for both engines together (all inflate bytes / time until the last block is inflated) is reported beside it.  `traffic`
    # is the kernel's measured DRAM bytes per algorithmic byte (one ncu --set full capture, profiles/inflate_traffic.json)
    # times its algorithmic bytes.
    sm_ms = sum(sm_inflate_ms) / max(len(sm_inflate_ms), 1)
    sm_share = stage_t.get("sm_inflate_share", 0.0)
    sm_bytes = alg_bytes * sm_share
    achieved = sm_bytes / (sm_ms * 1e-3) / 1e9 if sm_ms > 0 else 0.0
    traffic = None
    tp = os.path.join(ROOT, "profiles", "inflate_traffic.json")
    if os.path.exists(tp):
        try:
            ratio = json.load(open(tp)).get("dram_bytes_per_algorithmic_byte")
            traffic = None if ratio is None else ratio * sm_bytes
        except Exception:
            traffic = None
    # ---------------- CPU baseline on a bounded sample (rank 0, N = 1 only)
    cpu = None
    if world == 1:
        try:
            oracle = build.build_oracle()
            sm = gen_cached(os.path.join(a.workdir, "cpu_sample"), genome=a.genome, introns=a.introns, pairs=a.cpu_sample_pairs, readlen=a.readlen, seed=2, sorted_bam=False, chunks=16)
            sec = oracle_run(oracle, sm, sm["collated"], os.path.join(a.workdir, "cpu_sample", "out"))
            cpu = {"value": round((sm["pairs"] + sm["singles"]) / sec / 1e6, 4), "unit": UNIT, "cores": 1, "kind": "port",
                   "sample": f"{a.cpu_sample_pairs / 1e6:g}M-pair collated BAM of the same workload shape, one process, gzip inflate + reference load included ({sec:.1f} s)"}
        except Exception as e:  # the baseline is reporting only; never let it kill the bench line
            cpu = {"value": None, "unit": UNIT, "cores": 1, "kind": "port", "sample": f"failed: {e}"}
    line = {"metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(dt / a.steps * 1e3, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8 bytes -> int32/u32 counts, f64 ratios", "data": "synthetic",
            "config": {"workload": workload_name(a, total_pairs), "fragments": n_units, "records": n_records, "compressed_bytes": comp_bytes,
                       "inflated_bytes": infl_bytes, "l2": "inputs larger than L2 (no flush needed)", "sharding": "one BGZF byte range (genomic region) of the BAM per rank, resident in that rank's HBM; halo exchange + NCCL table merge inside every step" if world > 1 else "single GPU",
                       "parity_nccl": parity_nccl, "parity_nccl_what": "configs[0] (1M pairs, chr22-like) through run_sharded over this NCCL group vs the CPU restatement, six files byte for byte" if world > 1 else None,
                       "parity_nccl_sha256": parity_sha or None, "parity_nccl_foreign_depth_entries": parity_foreign, "output_sha256": sha_dir(outdir),
                       "merge_phase_ms": (merge_phases if world > 1 else None), "generation_s": round(gen_s, 1), "ref_load_s": round(ref_s, 2), "pairs_counted": stats["pairs"], "checks": "ok" if all(checks.values()) else {k: v for k, v in checks.items() if not v}, "stage_ms": {k: round(v, 3) for k, v in stage_t.items()}},
            "e2e": {"value": round(e2e_value, 3), "unit": UNIT, "h2d_bytes_per_step": int(comp_bytes), "d2h_bytes_per_step": int(d2h), "ms_per_step": round(dte / a.steps * 1e3, 3), "text_bytes_written": int(out_bytes), "stage_ms": {k: round(v, 3) for k, v in e2e_stage.items()}},
            "e2e_file": e2e_file,
            "gpu_launches": launches,
            "roofline": {"bound": "hbm", "kernel": "inflate_tok_kernel (two-phase DEFLATE: Huffman decode per lane, LZ77 resolve per warp): %.0f %% of the inflate bytes in %.0f persistent launch(es) per step; the hardware decompression engine inflates the other %.0f %% concurrently" % (sm_share * 100, inflate_launches / max(a.steps, 1), 100 - sm_share * 100),
                         "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 5), "traffic": traffic, "traffic_source": "profiles/inflate_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum per algorithmic byte from one ncu --set full capture of this very launch (tools/ncu_inflate_benched.sh: the first SM-inflate launch of `python bench.py` on this workload) x this launch's algorithmic bytes", "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": int(sm_bytes), "launches_per_step": inflate_launches / max(a.steps, 1), "avg_launch_ms": round(sm_ms, 3),
                         "both_engines": {"achieved": round(stage_achieved, 2), "frac": round(stage_achieved / peak, 5), "algorithmic_bytes_per_step": int(alg_bytes), "ms_until_all_inflated": round(avg_inflate_ms, 3)}},
            "cpu_baseline": cpu, "clocks": clocks}
    emit(line)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


def numa_bind(local):
    """Pins this process to the CPUs of the NUMA node its GPU hangs off (first-touch then places the page-locked buffer there)."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bus = "%04x:%02x:%02x.0" % (getattr(p, "pci_domain_id", 0), p.pci_bus_id, p.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        if node < 0:
            return
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            print(f"[bench] rank on GPU {local} ({bus}) bound to NUMA node {node}: {len(cpus)} CPUs", file=sys.stderr)
    except Exception as e:   # topology files missing in a container, attribute names differ between torch versions: run unbound
        print(f"[bench] --numa-bind ignored: {e}", file=sys.stderr)


_RESULT_FD = None


def emit(line):
    """The ONE line of stdout.  Everything else this process and its libraries print (NCCL's version banner, build logs)
    goes to stderr: file descriptor 1 is pointed at stderr for the whole run and the original stdout is kept for this call."""
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


if __name__ == "__main__":
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
