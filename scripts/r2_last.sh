#!/bin/bash
set -u
out=gpurun_out/r2_last
mkdir -p $out
timeout 200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "fused or weight" > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 $out/pytest.log
CUDA_VISIBLE_DEVICES=0 timeout 100 python scripts/oracle_errors.py 2>&1 | grep -v Warn | tee $out/oracle_errors.txt
