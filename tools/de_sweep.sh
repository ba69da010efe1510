#!/bin/bash
# developer helper: full-size bench at several decompression-engine shares
for p in "$@"; do
  python bench.py --steps 2 --warmup 1 --de-percent $p --cpu-sample-pairs 100000 2>/dev/null | tail -1 > /tmp/de_sweep.json
  python - "$p" <<'PY'
import json, sys
d = json.load(open("/tmp/de_sweep.json"))
print("de_percent", sys.argv[1], "value", d["value"], "e2e", d["e2e"]["value"], "inflate_ms", d["config"]["stage_ms"]["inflate_kernel_ms"], "e2e_inflate_wall_ms", d["e2e"]["stage_ms"]["inflate_wall_ms"], d["config"]["checks"])
PY
done
