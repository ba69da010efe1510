#!/bin/bash
# Per-instruction stall reasons of the two-phase inflate kernel (run on the GPU box):
#   gpurun --timeout 900 -- 'bash tools/ncu_tok_source.sh [variant]'
# One full grid of blocks (5M pairs ~ 113k BGZF blocks), SM kernel only.  The report stays in /tmp on the box (tens of MB);
# only the CSV pages come back: gpurun_out/tok_source.csv (page "source": SASS with sampling data per instruction) and tok_raw.csv.
set -u
V=${1:-12}; OUT=gpurun_out; mkdir -p $OUT
CMD="python tools/quick_bench.py --pairs 5000000 --genome grch38 --introns 200000 --readlen 150 --variants $V --reps 1 --out /tmp/qb_src"
$CMD > $OUT/tok_source_plain.log 2>&1 || exit 1
timeout 600 ncu --set full --import-source on --clock-control none -k regex:inflate_tok -c 1 -o /tmp/tok_src $CMD > $OUT/tok_source_ncu.log 2>&1
ncu -i /tmp/tok_src.ncu-rep --page raw --csv > $OUT/tok_raw.csv 2>/dev/null
ncu -i /tmp/tok_src.ncu-rep --page source --csv > $OUT/tok_source.csv 2>/dev/null
ls -la /tmp/tok_src.ncu-rep $OUT/tok_source.csv $OUT/tok_raw.csv
head -c 600 $OUT/tok_source.csv
