P='import sys,json; l=json.loads(sys.stdin.read()); print(sys.argv[1], l["config"]["workload"], "ms", round(l["ms_per_step"],3), "TF", round(l["roofline"]["achieved"],1), "frac", round(l["roofline"]["frac"],3), l["clocks"]["sm_mhz"], l["clocks"]["reasons"])'
for i in 1 2; do
for lib in libhop_sm100.so libhop_sm100_st9.so; do
HOP_SM100_LIB=$PWD/hop-vae_b200/$lib timeout 200 python bench.py --workload sweep_p64k_d512 --no-cpu-baseline --steps 20 2>&1 | grep '^{' | python -c "$P" $lib
done
done
