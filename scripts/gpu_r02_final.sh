#!/bin/bash
# round 2, final single-GPU collection on the committed tree: bench line, backward, sweep, model configs, launch lists
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02z_bench.json 2> gpurun_out/r02z_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02z_bench_reference.json 2> gpurun_out/r02z_bench_reference.err; echo "ref rc=$?"
timeout 600 python scripts/bench_backward.py > gpurun_out/r02z_backward.jsonl 2> gpurun_out/r02z_backward.err; echo "bwd rc=$?"
timeout 1500 python scripts/sweep_config5.py > gpurun_out/r02z_sweep_config5.jsonl 2> gpurun_out/r02z_sweep.err; echo "sweep rc=$?"
timeout 900 python scripts/bench_model_configs.py > gpurun_out/r02z_model_configs.jsonl 2> gpurun_out/r02z_model_configs.err; echo "configs rc=$?"
BWD="python scripts/bench_backward.py --points=16384,65536,256"
timeout 300 $BWD > gpurun_out/r02z_bwd_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02z_bwd_launches.csv $BWD > gpurun_out/r02z_bwd_launches_ncu.log 2>&1
echo "ncu rc=$?"
