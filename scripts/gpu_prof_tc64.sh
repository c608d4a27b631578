# launch list + one ncu --set full capture of the 64-rows-per-CTA pair kernel: bash scripts/gpu_prof_tc64.sh <workload> <tag>
W=${1:-sweep_p64k_d512}; T=${2:-tc64_d512}
mkdir -p gpurun_out
CMD="python bench.py --workload $W --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_$T.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches_$T.csv $CMD > gpurun_out/ncu_${T}_1.log 2>&1
$CMD > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:hop_fwd_tc -s 6 -c 1 -o gpurun_out/prof_$T $CMD > gpurun_out/ncu_${T}_2.log 2>&1
tail -2 gpurun_out/ncu_${T}_2.log
ncu -i gpurun_out/prof_$T.ncu-rep --page raw --csv > gpurun_out/prof_${T}_raw.csv 2>/dev/null
grep '^{' gpurun_out/plain_$T.log | cut -c1-200
