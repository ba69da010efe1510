#!/bin/bash
# Developer helper for the next round: A/B the prepared inflate variants (0 = default, 11 = byte distance LUT, 10 = the same
# with six CTAs per SM) on the SM kernel alone, then the whole resident step with the faster one.
#   gpurun --timeout 900 -- 'bash tools/next_round_ab.sh'
set -e
python tools/quick_bench.py --pairs 20000000 --readlen 150 --genome grch38 --introns 200000 --variants 0,11,10 --reps 2 2>&1 | tail -4
for v in 0 10; do
  python bench.py --steps 3 --warmup 2 --inflate-variant $v --cpu-sample-pairs 100000 2>/dev/null | python -c "
import json, sys
d = json.loads(sys.stdin.readline())
print('variant', $v, 'value', d['value'], 'e2e', d['e2e']['value'], 'sm inflate ms', d['config']['stage_ms']['sm_inflate_ms'], 'share', d['config']['stage_ms']['sm_inflate_share'], d['config']['checks'])"
done
