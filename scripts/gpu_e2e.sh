mkdir -p gpurun_out
show() { python - "$1" "$2" <<'PY'
import json,sys
for ln in open(sys.argv[1]):
    if ln.startswith("{"):
        j=json.loads(ln); print(sys.argv[2], "value %.4g ms/step %.3f kernel_ms %.3f frac %.3f e2e %.4g numa %s"%(j["value"], j["ms_per_step"], j["roofline"]["kernel_ms"], j["roofline"]["frac"], j["e2e"]["value"], j["config"].get("host_numa_cpus")))
PY
}
for c in default 1,3,3,1 1,2,2,2,1 1,2,3,2 4; do
  if [ $c = default ]; then X=""; else X="--e2e-chunks $c"; fi
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline $X > gpurun_out/e2e_$c.log 2>&1; show gpurun_out/e2e_$c.log "chunks $c"
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/e2e_g2.log 2>&1; show gpurun_out/e2e_g2.log "g2 batch"
tail -3 gpurun_out/e2e_g2.log | cut -c1-300
