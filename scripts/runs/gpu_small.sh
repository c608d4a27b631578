mkdir -p gpurun_out
python scripts/prof_small.py fwd 200 > gpurun_out/small_fwd_host.log 2>&1
python scripts/prof_small.py fwdbwd 100 > gpurun_out/small_fwdbwd_host.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_fwd_launches.csv python scripts/prof_small.py fwd 2 > gpurun_out/ncu_s1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_fwdbwd_launches.csv python scripts/prof_small.py fwdbwd 2 > gpurun_out/ncu_s2.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/small_fwdbwd_launches_ffhq64.csv python scripts/prof_small.py fwdbwd 2 FFHQ_64 > gpurun_out/ncu_s3.log 2>&1
python scripts/bench_model_configs.py > gpurun_out/model_configs.jsonl 2> gpurun_out/model_configs.err
tail -3 gpurun_out/model_configs.jsonl | cut -c1-700
