# usage: bash scripts/gpu_multi.sh <ngpus>
G=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q -s > gpurun_out/sharded_tests_g$G.log 2>&1; echo "pytest rc=$?" >> gpurun_out/sharded_tests_g$G.log
grep -E 'rel-L2|passed|failed|skipped|rc=|Error' gpurun_out/sharded_tests_g$G.log | tail -20
for mode in batch sharded; do
  for wl in sweep_p64k_d256 sweep_p256k_d256; do
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $G --steps 10 --warmup 3 --mode $mode --workload $wl > gpurun_out/bench_${mode}_${wl}_g$G.log 2>&1; echo "rc=$?" >> gpurun_out/bench_${mode}_${wl}_g$G.log
    grep -E '^\{|rc=[1-9]' gpurun_out/bench_${mode}_${wl}_g$G.log | python -c "
import sys, json
for ln in sys.stdin:
    if ln.startswith('{'):
        j=json.loads(ln); print('$mode $wl g$G: value %.4g ms/step %.3f kernel_ms %.3f frac %.3f e2e %.4g'%(j['value'], j['ms_per_step'], j['roofline']['kernel_ms'], j['roofline']['frac'], j['e2e']['value']))
    else: print('$mode $wl', ln)
"
  done
done
