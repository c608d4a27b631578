mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_v5.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_tc_v5.csv $CMD > gpurun_out/ncu_v5_1.log 2>&1
$CMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:hop_fwd_tc -s 6 -c 1 -o gpurun_out/prof_tc_v5 $CMD > gpurun_out/ncu_v5_2.log 2>&1
tail -2 gpurun_out/ncu_v5_2.log
grep '^{' gpurun_out/plain_v5.log | cut -c1-200
