G=${1:-2}
mkdir -p gpurun_out
for b in 0 148 74; do
  if [ $b = 0 ]; then unset HOP_PEER_BLOCKS; else export HOP_PEER_BLOCKS=$b; fi
  echo "HOP_PEER_BLOCKS=$b"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 29512 scripts/bench_peer_merge.py 2>&1 | grep -E '^\{|Error|error' | tail -3
done
