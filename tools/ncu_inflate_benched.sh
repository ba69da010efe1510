#!/bin/bash
# DRAM traffic of the dominant kernel on the launch bench.py times (run on the GPU box):
#   gpurun --timeout 900 -- 'bash tools/ncu_inflate_benched.sh'
# The first inflate_tok_kernel launch of `python bench.py` is the resident pipeline's SM share of the 50M-pair file.  Writes
# gpurun_out/inflate_benched_raw.csv (ncu raw page, one launch) and gpurun_out/inflate_traffic.json; copy the latter to profiles/.
set -u
OUT=gpurun_out; mkdir -p $OUT
python bench.py --steps 1 --warmup 1 --cpu-sample-pairs 100000 > $OUT/inflate_benched_plain.json 2> /dev/null || exit 1
timeout 800 ncu --set full --clock-control none -k regex:inflate_tok -c 1 -o /tmp/inflate_benched python bench.py --steps 1 --warmup 1 --cpu-sample-pairs 100000 > $OUT/inflate_benched_ncu.log 2>&1
ncu -i /tmp/inflate_benched.ncu-rep --page raw --csv > $OUT/inflate_benched_raw.csv 2>/dev/null
python tools/ncu_digest.py $OUT/inflate_benched_raw.csv $OUT/inflate_benched_digest.json > /dev/null
python - <<'PY'
import json
d = json.load(open("gpurun_out/inflate_benched_digest.json"))["launches"][0]
b = json.load(open("gpurun_out/inflate_benched_plain.json"))
alg = b["roofline"]["algorithmic_bytes_per_launch"] / max(b["roofline"]["launches_per_step"], 1)
out = {"kernel": d["kernel"], "capture": "ncu --set full --clock-control none -k regex:inflate_tok -c 1 python bench.py --steps 1 --warmup 1 (first SM-inflate launch of the resident pipeline: the launch bench.py times)",
       "algorithmic_bytes": int(alg), "dram_bytes_read": int(d["dram_read_GB"] * 1e9), "dram_bytes_write": int(d["dram_write_GB"] * 1e9),
       "dram_bytes_per_algorithmic_byte": round((d["dram_read_GB"] + d["dram_write_GB"]) * 1e9 / alg, 3), "gpu_time_ms_under_ncu": d["duration_ms"],
       "issue_active_pct": d["issue_active_pct"], "active_threads_per_instruction": d["threads_per_instruction"], "l1_hit_pct": d["l1_hit_pct"], "l2_hit_pct": d["l2_hit_pct"],
       "global_load_sectors_per_request": d["global_load_sectors_per_request"], "registers": d["registers"], "warp_instructions": d["warp_instructions"]}
json.dump(out, open("gpurun_out/inflate_traffic.json", "w"), indent=1)
print(json.dumps(out))
PY
