mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/t2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t2.log
python bench.py --workload sweep_p64k_d64 --path fp32 --steps 5 --warmup 3 > gpurun_out/bench_simt.log 2>&1; echo "rc=$?" >> gpurun_out/bench_simt.log
CMD="python bench.py --workload sweep_p64k_d64 --path fp32 --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_simt.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:hop_fwd_simt -s 3 -c 1 -o gpurun_out/prof_simt $CMD > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/t2.log; cat gpurun_out/bench_simt.log
