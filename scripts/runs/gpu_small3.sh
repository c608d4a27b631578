mkdir -p gpurun_out
python scripts/prof_small.py fwd 2 FFHQ_64 > /dev/null 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_fwd_launches_ffhq64.csv python scripts/prof_small.py fwd 2 FFHQ_64 > gpurun_out/ncu_s4.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:hop_fwd_simt -s 9 -c 3 -o gpurun_out/prof_simt_fwd python scripts/prof_small.py fwd 2 FFHQ_64 > gpurun_out/ncu_s5.log 2>&1
tail -2 gpurun_out/ncu_s5.log
