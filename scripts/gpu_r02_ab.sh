#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_parity_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q -s > gpurun_out/r02m_tests.txt 2>&1; echo "pytest rc=$?"
grep -E "passed|failed|block check|rows re-scored|Error" gpurun_out/r02m_tests.txt | tail -16
timeout 1500 python scripts/sweep_config5.py "--points=65536,256,0;65536,64,1;65536,128,3;262144,256,3;1048576,256,3;65536,256,3" > gpurun_out/r02m_sweep.jsonl 2> gpurun_out/r02m_sweep.err; echo "rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/r02m_sweep.jsonl"):
    d=json.loads(l); print(d["patterns"],d["dim"],d["update_steps"],"ms %.3f frac %.3f rel %.2e mism %d/%d flagged %d rescored %d"%(d["ms"],d["frac_of_measured_peak"],d["rel_l2"],d["argmax_mismatch"],d["argmax_decided_rows"],d["rows_failed_certificate"],d["rows_rescored_fp32"]), d.get("ms_without_exact_index"))
PY
tail -3 gpurun_out/r02m_sweep.err
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-sustained --no-backward 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench value %.2fM frac %.3f kernel_ms %.4f'%(d['value']/1e6,d['roofline']['frac'],d['roofline']['kernel_ms']))"
