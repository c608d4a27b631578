mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q -s -k "ddp" > gpurun_out/ddp_test.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ddp_test.log
grep -E 'rank 0|  d |passed|failed|rc=|Error' gpurun_out/ddp_test.log | tail -12
timeout 300 python scripts/bench_ffhq64_ddp.py 20 2>&1 | grep -E '^\{|Error' | tee gpurun_out/ffhq64_ddp_g1.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 scripts/bench_ffhq64_ddp.py 20 2>&1 | grep -E '^\{|Error' | tee gpurun_out/ffhq64_ddp_g2.json
