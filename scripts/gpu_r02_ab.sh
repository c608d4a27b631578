#!/bin/bash
P="--points=16384,65536,256"
for t in default nofence bufs2 default nofence bufs2; do
  if [ $t = default ]; then L=""; else L=$PWD/hop-vae_b200/libhop_sm100_$t.so; fi
  echo -n "$t "; HOP_SM100_LIB=$L timeout 300 python scripts/bench_backward.py $P 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_backward'])"
done
