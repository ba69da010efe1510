#!/bin/bash
# Per-kernel ncu captures of the counting side (run on the GPU box; writes small CSVs into gpurun_out/, never the .ncu-rep itself):
#   gpurun --timeout 900 -- 'bash tools/ncu_kernels.sh'
# One launch of every kernel over the whole 20M-pair file (stage-by-stage path, SM inflate only), `--set full`, then the
# "details" and "raw" pages as CSV.
set -u
OUT=gpurun_out; mkdir -p $OUT
K='regex:process_pairs|pair_records|crc_check|walk_count|frame_guess|intron_stats|scan_apply|scan_tile|max_depth|gather_records|process_singles|walk_list|ht_compact'
timeout 600 ncu --set full --clock-control none -k "$K" -o /tmp/r2_kernels python tools/quick_bench.py --pairs 20000000 --readlen 150 --genome grch38 --introns 200000 --variants 12 --reps 1 > $OUT/ncu_k.log 2>&1
tail -c 200 $OUT/ncu_k.log
ncu -i /tmp/r2_kernels.ncu-rep --page details --csv > $OUT/r2_ncu_kernels_details.csv 2>/dev/null
ncu -i /tmp/r2_kernels.ncu-rep --page raw --csv > $OUT/r2_ncu_kernels_raw.csv 2>/dev/null
ls -la $OUT/r2_ncu_kernels_*.csv
