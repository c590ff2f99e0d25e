#!/bin/bash
# Round-2 single-GPU check: GPU tests, flagship bench (single launch vs bucketed K1 with short-lived blocks), probes.
set -u
out=gpurun_out/r2_n1
mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/summary.txt
tail -n 12 $out/pytest.log
b="timeout 300 python bench.py --steps 200 --warmup 20 --no-e2e"
timeout 300 python bench.py --steps 200 --warmup 20 > $out/bench_default.json 2> $out/bench_default.err
export MTB_OVERLAP_N1=1
$b > $out/bench_ov.json 2>&1
MTB_K1_BLOCK_A=128 $b > $out/bench_ov_b128.json 2>&1
MTB_K1_BLOCK_A=32 $b > $out/bench_ov_b32.json 2>&1
MTB_COMM_PRIO=-1 $b > $out/bench_ov_prio_hi.json 2>&1
MTB_K1_GRID_A=148 MTB_K1_BLOCK_A=128 $b > $out/bench_ov_persist148.json 2>&1
MTB_K1_PROBE=lean:60 MTB_K1_GRID_A=148 MTB_K1_BLOCK_A=128 $b > $out/bench_probe_lean.json 2>&1
MTB_K1_PROBE=fat:60 MTB_K1_GRID_A=148 MTB_K1_BLOCK_A=128 $b > $out/bench_probe_fat.json 2>&1
MTB_K1_PROBE=fat:60 MTB_K1_GRID_A=296 MTB_K1_BLOCK_A=128 $b > $out/bench_probe_fat296.json 2>&1
MTB_K1_PROBE=lean:1 MTB_K1_GRID_A=1 MTB_K1_BLOCK_A=32 $b > $out/bench_probe_forkjoin.json 2>&1
unset MTB_OVERLAP_N1
$b --dtype fp32 --batch 100 > $out/bench_fp32_b100.json 2>&1
python scripts/benchsum.py $out/bench_*.json | tee -a $out/summary.txt
