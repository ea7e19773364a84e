#!/bin/bash
# ncu DRAM-traffic / L2-hit experiments on the dW GEMM of a 16384-wide layer (results in gpurun_out/dw_l2.txt).
# Each line: configuration, then for the one captured launch: duration, DRAM bytes read, L2 sector hit rate, SM clock.
out=gpurun_out/dw_l2.txt; : > $out
run() { # label, env..., uses $W $R
  label=$1; shift
  env "$@" REPS=2 timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct,sm__cycles_elapsed.avg.per_second \
     --clock-control none -k regex:umma_pair -s 1 -c 1 --csv python tools/profile_wide.py $W $R 2048 > gpurun_out/_dw_l2_one.csv 2>&1
  python - "$label" >> $out <<'PY'
import csv, sys
vals = {}
for row in csv.reader(open("gpurun_out/_dw_l2_one.csv", errors="replace")):
    if len(row) > 3 and row[-3] in ("gpu__time_duration.sum", "dram__bytes_read.sum", "lts__t_sector_hit_rate.pct",
                                     "sm__cycles_elapsed.avg.per_second"):
        vals[row[-3]] = f"{row[-1]} {row[-2]}"
print(sys.argv[1], "|", "; ".join(f"{k}={v}" for k, v in vals.items()) or "no data")
PY
}
W=16384; R=65536
run "one piece, group_m 8 (default), 74 pairs" DW_CHUNK=65536
run "one piece, group_m 4" DW_CHUNK=65536 DDNN_DBG_GROUP_M=4
run "one piece, group_m 16" DW_CHUNK=65536 DDNN_DBG_GROUP_M=16
run "one piece, group_m 8, 72 pairs (8 x 9 waves)" DW_CHUNK=65536 DDNN_DBG_PAIRS=72
run "one piece, group_m 8, 64 pairs (8 x 8 waves)" DW_CHUNK=65536 DDNN_DBG_PAIRS=64
run "chunk of 16384 rows" DW_CHUNK=16384
run "chunk of 8192 rows" DW_CHUNK=8192
run "chunk of 2048 rows" DW_CHUNK=2048
W=16640; R=32768
run "width 16640 (row pitch 33280 B, not a power of two), 32768 rows" DW_CHUNK=32768
W=16384; R=32768
run "width 16384, 32768 rows" DW_CHUNK=32768
rm -f gpurun_out/_dw_l2_one.csv
cat $out
