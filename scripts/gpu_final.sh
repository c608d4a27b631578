mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/final_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final_tests.log
tail -4 gpurun_out/final_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/final_smoke.log
tail -3 gpurun_out/final_smoke.log
timeout 600 python bench.py --impl reference > gpurun_out/final_bench_ref.log 2>&1; echo "rc=$?" >> gpurun_out/final_bench_ref.log
tail -2 gpurun_out/final_bench_ref.log
for i in 1 2; do
timeout 600 python bench.py > gpurun_out/final_bench_$i.log 2>&1; echo "rc=$?" >> gpurun_out/final_bench_$i.log
tail -2 gpurun_out/final_bench_$i.log
done
