#!/bin/bash
set -u
out=gpurun_out/r2_n1h
mkdir -p $out
timeout 400 python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 6 $out/pytest.log
timeout 150 python bench.py --steps 300 --warmup 20 > $out/bench_default.json 2> $out/bench_default.err
python scripts/benchsum.py $out/bench_*.json
