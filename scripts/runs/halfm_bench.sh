mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/halfm_all_tests.log 2>&1; echo "rc=$?" >> gpurun_out/halfm_all_tests.log
tail -3 gpurun_out/halfm_all_tests.log
for w in sweep_p64k_d512 sweep_p64k_d1024 sweep_p256k_d512 sweep_p64k_d512_t1; do
timeout 200 python bench.py --workload $w --no-cpu-baseline > gpurun_out/halfm_bench_$w.log 2>&1; echo "rc=$?" >> gpurun_out/halfm_bench_$w.log
grep '^{' gpurun_out/halfm_bench_$w.log | python -c "import sys,json; l=json.loads(sys.stdin.read()); print(l['config']['workload'], 'ms', round(l['ms_per_step'],3), 'value', round(l['value']), 'e2e', round(l['e2e']['value']), 'TF', round(l['roofline']['achieved'],1), 'frac', round(l['roofline']['frac'],3), l['clocks'])"
done
HOP_TC_HALFM=0 timeout 200 python bench.py --workload sweep_p64k_d512 --no-cpu-baseline 2>&1 | grep '^{' | python -c "import sys,json; l=json.loads(sys.stdin.read()); print('HALFM=0', l['config']['workload'], 'ms', round(l['ms_per_step'],3), 'TF', round(l['roofline']['achieved'],1), 'frac', round(l['roofline']['frac'],3))"
