mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/d1024_all_tests.log 2>&1; echo "rc=$?" >> gpurun_out/d1024_all_tests.log
tail -4 gpurun_out/d1024_all_tests.log
for w in sweep_p64k_d1024 sweep_p64k_d1024_t1 sweep_p64k_d512; do
timeout 200 python bench.py --workload $w --no-cpu-baseline > gpurun_out/d1024_bench_$w.log 2>&1; echo "rc=$?" >> gpurun_out/d1024_bench_$w.log
tail -2 gpurun_out/d1024_bench_$w.log | cut -c1-1500
done
