#!/bin/bash
set -u
out=gpurun_out/r2_n1d
mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_fp32.py -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/summary.txt
tail -n 5 $out/pytest.log
b="timeout 100 python bench.py --steps 200 --warmup 20 --no-e2e"
$b --dtype fp32 --batch 100 > $out/bench_fp32_b100.json 2>&1
timeout 100 python bench.py --steps 30 --warmup 5 --model wide --optimizer adam --batch 256 --rows 4096 --no-e2e > $out/bench_wide_adam.json 2>&1
python scripts/benchsum.py $out/bench_*.json | tee -a $out/summary.txt
cmd="python bench.py --steps 3 --warmup 3 --no-e2e --no-graph --dtype fp32 --batch 100"
timeout 100 $cmd > $out/plain.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file $out/launches_fp32.csv $cmd > $out/ncu.log 2>&1
python scripts/launches.py $out/launches_fp32.csv conv1_fwd 2>&1 | tail -n 24 | tee $out/launches_fp32.txt
cmd="python bench.py --steps 3 --warmup 3 --no-e2e --no-graph --model wide --optimizer adam --batch 256 --rows 4096"
timeout 100 $cmd > $out/plain2.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file $out/launches_wide.csv $cmd > $out/ncu2.log 2>&1
python scripts/launches.py $out/launches_wide.csv conv1_fwd 2>&1 | tail -n 24 | tee $out/launches_wide.txt
