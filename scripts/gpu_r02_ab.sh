#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_backward_tc_gpu.py tests/test_parity_gpu.py -m gpu -x -q -k "backward or train or grad" > gpurun_out/r02n_tests.txt 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r02n_tests.txt
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 600 python scripts/bench_backward.py > gpurun_out/r02n_backward.jsonl 2>/dev/null; python -c "
import json
for l in open('gpurun_out/r02n_backward.jsonl'):
    b=json.loads(l); print(b['rows'],b['patterns'],b['dim'],'fwd+prep %.2f bwd %.2f frac %.3f'%(b['ms_forward_incl_prep'],b['ms_backward'],b['bwd_frac_of_measured_peak']))"
BWD="python scripts/bench_backward.py --points=16384,65536,256"
timeout 300 $BWD > gpurun_out/r02n_bwd_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02n_bwd_launches.csv $BWD > gpurun_out/r02n_bwd_launches_ncu.log 2>&1
echo "ncu rc=$?"
