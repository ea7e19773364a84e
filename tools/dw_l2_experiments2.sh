#!/bin/bash
# ncu: DRAM traffic of ONE 8192-row dW chunk (and of one forward micro-batch) of a 16384-wide layer vs the
# rasterisation group of the CTA-pair kernel (results in gpurun_out/dw_l2_chunk.txt).
out=gpurun_out/dw_l2_chunk.txt; : > $out
run() {
  label=$1; shift
  env "$@" REPS=2 timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__cycles_elapsed.avg.per_second \
     --clock-control none -k regex:umma_pair -s 1 -c 1 --csv python tools/profile_wide.py 16384 65536 2048 > gpurun_out/_dw_l2_one.csv 2>&1
  python - "$label" >> $out <<'PY'
import csv, sys
vals = {}
for row in csv.reader(open("gpurun_out/_dw_l2_one.csv", errors="replace")):
    if len(row) > 3 and row[-3] in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
                                     "sm__cycles_elapsed.avg.per_second"):
        vals[row[-3]] = f"{row[-1]} {row[-2]}"
print(sys.argv[1], "|", "; ".join(f"{k}={v}" for k, v in vals.items()) or "no data")
PY
}
for gm in 2 4 8 16 32 64; do run "8192-row chunk, group_m $gm" DW_CHUNK=8192 DDNN_DBG_GROUP_M=$gm; done
for gm in 8 16 32; do run "16384-row chunk, group_m $gm" DW_CHUNK=16384 DDNN_DBG_GROUP_M=$gm; done
rm -f gpurun_out/_dw_l2_one.csv
cat $out
