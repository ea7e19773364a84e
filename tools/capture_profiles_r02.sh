# Round-2 ncu captures (run under gpurun on one GPU; every command below first runs without ncu).
set -x
M="gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum"
for wl in mnist4096 narrow65536; do
  timeout 120 python tools/profile_step.py $wl 3 > gpurun_out/plain_$wl.log 2>&1 && \
  timeout 300 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/r02_launches_${wl}_step.csv python tools/profile_step.py $wl 3 > gpurun_out/ncu_$wl.log 2>&1
done
timeout 300 python bench.py --steps 2 --warmup 1 --no-extra > gpurun_out/plain_bench.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2600 --csv --log-file gpurun_out/r02_launches_bench_wide8x16384.csv python bench.py --steps 2 --warmup 1 --no-extra > gpurun_out/ncu_bench.log 2>&1
timeout 380 ncu --set full --clock-control none --import-source on -k "regex:tf32_gemm" -s 2 -c 2 -o gpurun_out/prof_tf32_r02 -f python tools/profile_step.py mnist4096 2 > gpurun_out/ncu_out1.log 2>&1
ncu -i gpurun_out/prof_tf32_r02.ncu-rep --page raw --csv > gpurun_out/r02_prof_mnist_tf32_raw.csv 2>/dev/null
rm -f gpurun_out/prof_tf32_r02.ncu-rep
for wl in mnist4096 narrow65536; do python tools/launch_summary.py gpurun_out/r02_launches_${wl}_step.csv 12; done
