out=gpurun_out/dw_hints.txt; : > $out
run() {
  label=$1; shift
  env "$@" REPS=2 DW_CHUNK=8192 timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct,sm__cycles_elapsed.avg.per_second \
     --clock-control none -k regex:umma_pair -s 1 -c 1 --csv python tools/profile_wide.py 16384 65536 2048 > gpurun_out/_one.csv 2>&1
  python - "$label" >> $out <<'PY'
import csv, sys
vals = {}
for row in csv.reader(open("gpurun_out/_one.csv", errors="replace")):
    if len(row) > 3 and row[-3] in ("gpu__time_duration.sum", "dram__bytes_read.sum", "lts__t_sector_hit_rate.pct", "sm__cycles_elapsed.avg.per_second"):
        vals[row[-3]] = f"{row[-1]} {row[-2]}"
print(sys.argv[1], "|", "; ".join(f"{k}={v}" for k, v in vals.items()) or "no data")
PY
}
run "8192-row chunk, hints normal" X=1
run "A evict_last" DDNN_DBG_HINT_A=2
run "B evict_last" DDNN_DBG_HINT_B=2
run "A and B evict_last" DDNN_DBG_HINT_A=2 DDNN_DBG_HINT_B=2
run "A evict_last, B evict_first" DDNN_DBG_HINT_A=2 DDNN_DBG_HINT_B=1
run "A evict_first" DDNN_DBG_HINT_A=1
run "group_m 16, A evict_last" DDNN_DBG_GROUP_M=16 DDNN_DBG_HINT_A=2
rm -f gpurun_out/_one.csv; cat $out
