set -x
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
for wl in mnist4096 narrow65536; do
  timeout 120 python tools/profile_step.py $wl 3 > gpurun_out/plain_$wl.log 2>&1 && \
  timeout 300 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/r01d_launches_${wl}_step.csv python tools/profile_step.py $wl 3 > gpurun_out/ncu_$wl.log 2>&1
done
timeout 200 python tools/profile_step.py wide4x8192 2 8192 > gpurun_out/plain_wide.log 2>&1 && \
timeout 300 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/r01d_launches_wide4x8192_step.csv python tools/profile_step.py wide4x8192 2 8192 > gpurun_out/ncu_widel.log 2>&1
timeout 380 ncu --set full --clock-control none --import-source on -k "regex:tf32_gemm|tail2" -s 3 -c 3 -o gpurun_out/prof_tf32_d -f python tools/profile_step.py mnist4096 2 > gpurun_out/ncu_out1.log 2>&1
timeout 380 ncu --set full --clock-control none --import-source on -k "regex:outlayer_mma|umma" -s 6 -c 6 -o gpurun_out/prof_narrow_d -f python tools/profile_step.py narrow65536 2 > gpurun_out/ncu_out2.log 2>&1
timeout 300 python bench.py > gpurun_out/bench_default.log 2> gpurun_out/bench_default.err
tail -2 gpurun_out/bench_default.err
