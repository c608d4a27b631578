#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python scripts/sweep_config5.py "--points=65536,1024,0;262144,1024,0;65536,1024,1;65536,128,0;262144,128,0;65536,64,0;65536,128,3;65536,64,1" > gpurun_out/r02z_sweep_other_widths.jsonl 2> gpurun_out/r02z_sweep_other.err; echo "rc=$?"
cut -c1-260 gpurun_out/r02z_sweep_other_widths.jsonl; tail -3 gpurun_out/r02z_sweep_other.err
