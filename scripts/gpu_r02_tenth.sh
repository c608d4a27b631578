#!/bin/bash
# allocation-free entry point: its test alone in a fresh process, then the whole -m gpu suite (the file collects last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -s > gpurun_out/r02k_noalloc_alone.txt 2>&1; echo "alone rc=$?"
tail -4 gpurun_out/r02k_noalloc_alone.txt
timeout 2700 python -m pytest tests -m gpu -x -q -rs > gpurun_out/r02k_tests.txt 2>&1; echo "suite rc=$?"
tail -4 gpurun_out/r02k_tests.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02k_smoke.txt 2>&1; echo "smoke rc=$?"
