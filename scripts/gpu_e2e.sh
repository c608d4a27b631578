mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "forward_host" > gpurun_out/fh_test.log 2>&1; tail -3 gpurun_out/fh_test.log
for c in 1 2 4 8; do
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-chunks $c > gpurun_out/bench_e2e_c$c.log 2>&1
python - <<PY
import json
for ln in open("gpurun_out/bench_e2e_c$c.log"):
    if ln.startswith("{"):
        j=json.loads(ln); print("chunks $c: value %.4g ms/step %.3f e2e %.4g (%.3f ms)"%(j["value"], j["ms_per_step"], j["e2e"]["value"], 16384/j["e2e"]["value"]*1e3))
PY
done
