#!/bin/bash
set -u
out=gpurun_out/r2_n1i
mkdir -p $out
timeout 400 python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 5 $out/pytest.log
b="timeout 100 python bench.py --steps 300 --warmup 20 --no-e2e"
$b --batch 100 > $out/bench_bf16_b100.json 2>&1
MTB_OVERLAP_N1=1 $b --batch 100 > $out/bench_bf16_b100_ov.json 2>&1
$b --batch 100 --dtype fp32 > $out/bench_fp32_b100.json 2>&1
MTB_OVERLAP_N1=1 $b --batch 100 --dtype fp32 > $out/bench_fp32_b100_ov.json 2>&1
python scripts/benchsum.py $out/bench_*.json
python -c "import __graft_entry__ as g; g.smoke()"
