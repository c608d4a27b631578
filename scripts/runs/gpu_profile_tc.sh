# launch list + one full capture of the fused tensor-core forward at the default bench workload
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_tc.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:hop_fwd_tc -s 6 -c 1 -o gpurun_out/prof_tc $CMD > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
