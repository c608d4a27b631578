P='import sys,json; l=json.loads(sys.stdin.read()); print(sys.argv[1], l["config"]["workload"], "ms", round(l["ms_per_step"],3), "TF", round(l["roofline"]["achieved"],1), "frac", round(l["roofline"]["frac"],3), l["clocks"]["sm_mhz"], l["clocks"]["reasons"])'
HOP_TC64_TILE=256 timeout 100 python -m pytest tests/test_parity_gpu.py -x -q -k "d1024 and not 128_row" --timeout 60 2>&1 | tail -2
for t in 256 128 256 128; do
HOP_TC64_TILE=$t timeout 200 python bench.py --workload sweep_p64k_d1024 --no-cpu-baseline 2>&1 | grep '^{' | python -c "$P" tile$t
done
