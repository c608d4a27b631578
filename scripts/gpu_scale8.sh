# usage: bash scripts/gpu_scale8.sh <ngpus>: sharded parity at every world size <= G, then batch-parallel and sharded bench lines
G=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q -s > gpurun_out/sharded_tests_g$G.log 2>&1; echo "pytest rc=$?" >> gpurun_out/sharded_tests_g$G.log
grep -E 'passed|failed|skipped|rc=' gpurun_out/sharded_tests_g$G.log | tail -5
grep -E 'rank 0.*peer' gpurun_out/sharded_tests_g$G.log | tail -12
run() { # mode workload extra
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $G --steps 10 --warmup 3 --mode $1 --workload $2 $3 --no-cpu-baseline > gpurun_out/bench_$1_$2_g$G$4.log 2>&1; echo "rc=$?" >> gpurun_out/bench_$1_$2_g$G$4.log
  grep -E '^\{|rc=[1-9]' gpurun_out/bench_$1_$2_g$G$4.log | python -c "
import sys, json
for ln in sys.stdin:
    if ln.startswith('{'):
        j=json.loads(ln); print('$1 $2 $3 g$G: value %.4g ms/step %.3f kernel_ms %.3f e2e %.4g'%(j['value'], j['ms_per_step'], j['roofline']['kernel_ms'], j['e2e']['value']))
    else: print('$1 $2 $3', ln)
"
}
run batch sweep_p64k_d256 "" ""
run sharded sweep_p256k_d256 "--exchange peer" _peer
run sharded sweep_p256k_d256 "--exchange collective" _coll
run sharded sweep_p64k_d256 "--exchange peer" _peer
run sharded sweep_p64k_d256 "--exchange collective" _coll
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29512 scripts/bench_peer_merge.py 2>&1 | grep -E '^\{' | tail -1
