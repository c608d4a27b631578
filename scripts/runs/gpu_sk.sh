mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "tensor_core or full_size or folded or forward_host" 2>&1 | tail -4
for sk in 0 1; do
  for wl in sweep_p64k_d256; do
    HOP_TC_STREAMK=$sk timeout 300 python bench.py --workload $wl --path f16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/sk${sk}_$wl.log 2>&1
    python - "gpurun_out/sk${sk}_$wl.log" "streamk=$sk $wl" <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        j=json.loads(ln); print(sys.argv[2], "value %.4g ms/step %.3f kernel_ms %.3f TF %.1f frac %.3f e2e %.4g"%(j["value"], j["ms_per_step"], j["roofline"]["kernel_ms"], j["roofline"]["achieved"], j["roofline"]["frac"], j["e2e"]["value"]))
PY
    grep -E "Error|error" gpurun_out/sk${sk}_$wl.log | head -3
  done
done
