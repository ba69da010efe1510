#!/usr/bin/env python
"""Full-size byte parity of the CUDA path against the CPU restatement on BASELINE.json's configs[1..3].

Test infrastructure (it executes oracle/irfinder_cpu, which only tests/, smoke() and bench.py's CPU legs may do).  Two steps so
that the GPU is not held idle while the single-threaded oracle runs (~6 minutes per 50M-pair file):

    python tests/full_parity.py start  --configs C2,C3,C4 --workdir /tmp/irf_parity
        per config: generate the coordinate-sorted and the name-collated copy with the bench seeds, start the oracle on the
        collated copy in the background (one core each), run the sorted copy through irf_run_bam + irf_write_outputs on the
        GPU, digest IRF_T_DEPTH on the device, delete the sorted copy.
    python tests/full_parity.py finish --workdir /tmp/irf_parity --out profiles/parity_full_r2.json
        wait for the oracles, then compare: sha256 + length of each of the six files, the stdout counter block, and the
        per-base depth digest (n, sum, position-weighted sums mod 2^64, max — the oracle's --depth-digest).

`--scale 0.01` shrinks every config for a quick check of the tool itself.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

FILES = ["IRFinder-IR-nondir.txt", "IRFinder-IR-dir.txt", "IRFinder-JuncCount.txt", "IRFinder-SpansPoint.txt", "IRFinder-ChrCoverage.txt", "IRFinder-ROI.txt"]


def sha_file(p):
    h = hashlib.sha256()
    with open(p, "rb") as f:
        for blk in iter(lambda: f.read(1 << 22), b""):
            h.update(blk)
    return h.hexdigest(), os.path.getsize(p)


def device_depth_digest(core):
    """(n, sum, w1, w2, max) of the finished depth table, computed on the GPU from the library's own buffer (int64 arithmetic
    wraps mod 2^64 exactly like the oracle's unsigned sums)."""
    import torch
    from irfinder_b200.distributed import _wrap
    ptr, n, _ = core.device_table("diff")     # after irf_finalize the difference counters have been scanned into depth in place
    dev = torch.device("cuda", torch.cuda.current_device())
    v = _wrap(ptr, n, "<i4", dev)
    s = w1 = w2 = 0
    mx = 0
    M = (1 << 64) - 1
    step = 1 << 26
    for lo in range(0, n, step):
        x = v[lo:lo + step].to(torch.int64)
        i = torch.arange(lo + 1, lo + 1 + x.numel(), dtype=torch.int64, device=dev)
        s = (s + int(x.sum().item())) & M
        w1 = (w1 + int((x * i).sum().item())) & M
        w2 = (w2 + int((x * (i * i)).sum().item())) & M
        mx = max(mx, int(x.max().item()))
    return [n, s, w1, w2, mx]


def start(a):
    from irfinder_b200 import IRFinderCore, build, synth
    oracle = build.build_oracle()
    build.build_tools()
    os.makedirs(a.workdir, exist_ok=True)
    state = {"configs": {}, "scale": a.scale}
    core = None
    for cfg in a.configs.split(","):
        kw = dict(synth.CONFIGS[cfg])
        kw["pairs"] = max(1000, int(kw["pairs"] * a.scale))
        if a.scale < 1:
            kw["introns"] = max(100, int(kw["introns"] * max(a.scale, 0.05)))
            if kw.get("hot_introns"):
                kw["hot_introns"] = max(10, int(kw["hot_introns"] * max(a.scale, 0.05)))
        wd = os.path.join(a.workdir, cfg)
        t0 = time.perf_counter()
        meta = synth.generate(wd, chunks=256, **kw)
        gen_s = time.perf_counter() - t0
        odir = os.path.join(wd, "oracle")
        os.makedirs(odir, exist_ok=True)
        cmd = [oracle, odir, meta["cover"], meta["sj"], meta["rc"], meta["roi"], meta["collated"], "--depth-digest", os.path.join(odir, "depth.digest")]
        sh = " ".join("'%s'" % c for c in cmd) + " > '%s' 2> '%s'; echo $? > '%s'" % (os.path.join(odir, "stdout.txt"), os.path.join(odir, "stderr.txt"), os.path.join(odir, "rc"))
        t_or = time.time()
        subprocess.Popen(["sh", "-c", sh], start_new_session=True, stdin=subprocess.DEVNULL, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        # GPU side, through the reference-facing calls with default options
        if core is not None:
            core.close()
        core = IRFinderCore(0)
        core.load_ref(meta["cover"], meta["sj"], meta["rc"], meta["roi"])
        t1 = time.perf_counter()
        core.run_bam(meta["sorted"])
        gdir = os.path.join(wd, "gpu")
        core.write_outputs(gdir)
        gpu_s = time.perf_counter() - t1
        with open(os.path.join(gdir, "stdout.txt"), "w") as f:
            f.write(core.stdout_text)
        dig = device_depth_digest(core)
        with open(os.path.join(gdir, "depth.digest"), "w") as f:
            f.write(" ".join(str(x) for x in dig) + "\n")
        sizes = {"sorted_bytes": os.path.getsize(meta["sorted"]), "collated_bytes": os.path.getsize(meta["collated"])}
        os.remove(meta["sorted"])
        state["configs"][cfg] = dict(generator=kw, workdir=wd, generation_s=round(gen_s, 1), gpu_wall_s=round(gpu_s, 3), oracle_started=t_or, stats=core.stats(),
                                     directionality=core.directionality, **sizes)
        print(f"[full_parity] {cfg}: generated in {gen_s:.0f} s, GPU run_bam + write_outputs {gpu_s:.2f} s, oracle running", file=sys.stderr)
        with open(os.path.join(a.workdir, "state.json"), "w") as f:
            json.dump(state, f, indent=1)
    if core is not None:
        core.close()


def finish(a):
    state = json.load(open(os.path.join(a.workdir, "state.json")))
    out = {"what": "CUDA path (irf_run_bam + irf_write_outputs, default options, coordinate-sorted BAM) vs oracle/irfinder_cpu (name-collated copy of the same reads) at full size",
           "scale": state["scale"], "configs": {}}
    all_ok = True
    for cfg, st in state["configs"].items():
        odir, gdir = os.path.join(st["workdir"], "oracle"), os.path.join(st["workdir"], "gpu")
        deadline = time.time() + a.timeout
        while not os.path.exists(os.path.join(odir, "rc")) and time.time() < deadline:
            time.sleep(2)
        rec = dict(generator=st["generator"], generation_s=st["generation_s"], gpu_wall_s=st["gpu_wall_s"], sorted_bytes=st["sorted_bytes"], collated_bytes=st["collated_bytes"],
                   gpu_stats=st["stats"], directionality=st["directionality"])
        if not os.path.exists(os.path.join(odir, "rc")):
            rec["result"] = "oracle did not finish in time"
            all_ok = False
            out["configs"][cfg] = rec
            continue
        rc = int(open(os.path.join(odir, "rc")).read().strip() or 1)
        rec["oracle_rc"] = rc
        rec["oracle_wall_s"] = round(os.path.getmtime(os.path.join(odir, "rc")) - st["oracle_started"], 1)
        files, same = {}, rc == 0
        for name in FILES:
            po, pg = os.path.join(odir, name), os.path.join(gdir, name)
            eo, eg = os.path.exists(po), os.path.exists(pg)
            if not eo and not eg:
                continue
            so = sha_file(po) if eo else (None, None)
            sg = sha_file(pg) if eg else (None, None)
            files[name] = dict(oracle_sha256=so[0], oracle_bytes=so[1], gpu_sha256=sg[0], gpu_bytes=sg[1], identical=so == sg and eo and eg)
            same &= files[name]["identical"]
        rec["files"] = files
        so_txt, sg_txt = open(os.path.join(odir, "stdout.txt")).read(), open(os.path.join(gdir, "stdout.txt")).read()
        rec["stdout_identical"] = so_txt == sg_txt
        rec["stdout_gpu"] = sg_txt
        do = open(os.path.join(odir, "depth.digest")).read().split() if os.path.exists(os.path.join(odir, "depth.digest")) else None
        dg = open(os.path.join(gdir, "depth.digest")).read().split()
        rec["depth_digest"] = dict(fields="n sum w1 w2 max", oracle=do, gpu=dg, identical=do == dg)
        same &= rec["stdout_identical"] and do == dg
        rec["result"] = "identical" if same else "MISMATCH"
        all_ok &= same
        out["configs"][cfg] = rec
        print(f"[full_parity] {cfg}: {rec['result']} (oracle {rec['oracle_wall_s']} s)", file=sys.stderr)
        if a.cleanup:
            for f in ("collated.bam",):
                p = os.path.join(st["workdir"], f)
                if os.path.exists(p):
                    os.remove(p)
    out["all_identical"] = all_ok
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps({k: v["result"] for k, v in out["configs"].items()}))
    return 0 if all_ok else 1


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("cmd", choices=["start", "finish"])
    ap.add_argument("--configs", default="C2,C3,C4")
    ap.add_argument("--workdir", default="/tmp/irf_parity")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "parity_full.json"))
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--timeout", type=float, default=1500)
    ap.add_argument("--cleanup", type=int, default=1)
    args = ap.parse_args()
    sys.exit(start(args) or 0 if args.cmd == "start" else finish(args))
