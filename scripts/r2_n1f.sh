#!/bin/bash
set -u
out=gpurun_out/r2_n1g
mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_engine.py tests/test_gpu_kernels.py tests/test_gpu_fp32.py -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 $out/pytest.log
b="timeout 150 python bench.py --steps 300 --warmup 20 --no-e2e"
$b > $out/bench_default.json 2> $out/bench_default.err
$b --dtype fp32 --batch 100 > $out/bench_fp32_b100.json 2>&1
$b --batch 100 > $out/bench_bf16_b100.json 2>&1
python scripts/benchsum.py $out/bench_*.json
cmd="python bench.py --steps 3 --warmup 3 --no-e2e --no-graph"
timeout 100 $cmd > $out/plain.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file $out/launches.csv $cmd > $out/ncu.log 2>&1
python scripts/launches.py $out/launches.csv conv1_fwd 2>&1 | tail -n 16 | tee $out/launches.txt
