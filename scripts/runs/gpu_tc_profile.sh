mkdir -p gpurun_out
CMD="python bench.py --workload sweep_p64k_d256 --path f16 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_tc.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:hop_fwd_tc -s 4 -c 1 -o gpurun_out/prof_tc $CMD > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu_tests.log
tail -5 gpurun_out/gpu_tests.log
