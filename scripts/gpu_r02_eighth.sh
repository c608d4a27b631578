#!/bin/bash
mkdir -p gpurun_out
timeout 120 python scripts/probes/bwd_once.py --simt-tails
timeout 900 compute-sanitizer --tool initcheck --track-unused-memory no python scripts/probes/bwd_once.py --simt-tails > gpurun_out/r02i_initcheck.txt 2>&1; echo "initcheck rc=$?"
grep -c "Uninitialized" gpurun_out/r02i_initcheck.txt; grep -A12 "Uninitialized" gpurun_out/r02i_initcheck.txt | head -60
timeout 900 compute-sanitizer --tool racecheck python scripts/probes/bwd_once.py --simt-tails > gpurun_out/r02i_racecheck.txt 2>&1; echo "racecheck rc=$?"
grep -c "hazard" gpurun_out/r02i_racecheck.txt; grep -B2 -A10 "hazard" gpurun_out/r02i_racecheck.txt | head -80
tail -5 gpurun_out/r02i_racecheck.txt
