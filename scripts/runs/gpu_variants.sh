# usage: bash scripts/gpu_variants.sh "<tag> <tag> ..." "<workload> ..."   ("base" = the product library)
mkdir -p gpurun_out
for tag in $1; do
  if [ "$tag" = base ]; then unset HOP_SM100_LIB; else export HOP_SM100_LIB=$PWD/hop-vae_b200/libhop_sm100_$tag.so; fi
  timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -s -k "tensor_core" > gpurun_out/tc_tests_$tag.log 2>&1; echo "$tag pytest rc=$? worst rel-L2: $(grep -o 'rel-L2 [0-9.e-]*' gpurun_out/tc_tests_$tag.log | sort -g -k2 | tail -1)"
  for wl in $2; do
    timeout 300 python bench.py --workload $wl --path f16 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${tag}_$wl.log 2>&1
    python - <<PY
import json
for ln in open("gpurun_out/bench_${tag}_$wl.log"):
    if ln.startswith("{"):
        j=json.loads(ln); print("$tag $wl", "ms/step %.3f kernel_ms %.3f TF %.1f frac %.3f clocks %s %s"%(j["ms_per_step"], j["roofline"]["kernel_ms"], j["roofline"]["achieved"], j["roofline"]["frac"], j["clocks"]["sm_mhz"], j["clocks"]["reasons"]))
PY
  done
done
