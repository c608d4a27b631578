#!/bin/bash
# round 2, fourth GPU pass (one GPU): cost of the runner-up tracking (A/B against a build without it, same box), the
# scaled-residual GEMM preparation, certificate tests; then the config-5 sweep and one full ncu capture of the headline launch
mkdir -p gpurun_out
B="python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-sustained --no-parity --no-backward --no-sharded-record"
for i in 1 2; do
  HOP_SM100_LIB=$PWD/hop-vae_b200/libhop_sm100_notop2.so timeout 300 $B > gpurun_out/r02d_ab_notop2_$i.json 2>> gpurun_out/r02d_ab.err
  timeout 300 $B > gpurun_out/r02d_ab_top2_$i.json 2>> gpurun_out/r02d_ab.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r02d_ab_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["ms_per_step"], d["roofline"]["kernel_ms"], d["roofline"]["frac"])
    except Exception as e: print(f, "unreadable", e)
PY
timeout 1200 python -m pytest tests/test_parity_gpu.py -m gpu -q -s -k "preparation or certificate or diffuse or full_size or config" > gpurun_out/r02d_tests.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r02d_tests.txt
grep -E "passed|failed|rc=" gpurun_out/r02d_tests.txt | tail -3
timeout 1500 python scripts/sweep_config5.py > gpurun_out/r02d_sweep_config5.jsonl 2> gpurun_out/r02d_sweep.err; echo "sweep rc=$?"
timeout 300 $B > gpurun_out/r02d_full_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:hop_fwd_tc_kernel -s 1 -c 1 -o gpurun_out/r02d_fwd_full $B > gpurun_out/r02d_full_ncu.log 2>&1
echo "ncu fwd full rc=$?"
ls -la gpurun_out | tail -14
