mkdir -p gpurun_out
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final3_smoke.log 2>&1; echo "rc=$?" >> gpurun_out/final3_smoke.log
tail -3 gpurun_out/final3_smoke.log
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/final3_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final3_tests.log
tail -3 gpurun_out/final3_tests.log
timeout 240 python bench.py > gpurun_out/final3_bench.log 2>&1; echo "rc=$?" >> gpurun_out/final3_bench.log
tail -2 gpurun_out/final3_bench.log | cut -c1-400
for w in sweep_p64k_d512 sweep_p64k_d1024; do
timeout 200 python bench.py --workload $w --no-cpu-baseline > gpurun_out/final3_bench_$w.log 2>&1
grep '^{' gpurun_out/final3_bench_$w.log | python -c "import sys,json; l=json.loads(sys.stdin.read()); print(l['config']['workload'], 'ms', round(l['ms_per_step'],3), 'value', round(l['value']), 'e2e', round(l['e2e']['value']), 'frac', round(l['roofline']['frac'],3))"
done
