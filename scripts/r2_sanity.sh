#!/bin/bash
# Last check of a round on one GPU: the whole GPU test tier, the smoke step, and ncu --set full rows of the flagship
# step's three FC GEMMs (the two-CTAs-per-SM build among them).
set -u
out=gpurun_out/r2_sanity
mkdir -p $out
timeout 300 python -m pytest tests -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 5 $out/pytest.log
python -c "import __graft_entry__ as g; g.smoke()"
cmd="python bench.py --steps 2 --warmup 3 --no-e2e --no-graph"
timeout 100 $cmd > $out/plain.log 2>&1 && timeout 200 ncu --set full --clock-control none --import-source on -k "regex:gemm_bf16_kernel" -s 9 -c 3 -o $out/prof_gemm $cmd > $out/ncu_gemm.log 2>&1
echo "ncu rc=$?"
