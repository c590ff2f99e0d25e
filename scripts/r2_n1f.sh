#!/bin/bash
set -u
out=gpurun_out/r2_n1f
mkdir -p $out
timeout 120 python scripts/bench_gemm.py 2>&1 | tee $out/bench_gemm.txt
timeout 300 python -m pytest tests/test_gpu_engine.py tests/test_gpu_kernels.py -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 $out/pytest.log
timeout 150 python bench.py --steps 300 --warmup 20 > $out/bench_default.json 2> $out/bench_default.err
python scripts/benchsum.py $out/bench_*.json
timeout 150 python bench.py --reference-semantics --impl baseline --steps 200 --warmup 10 > $out/faithful_base_n1.json 2> $out/faithful_base_n1.err; grep "^{" $out/faithful_base_n1.json | cut -c1-400; tail -n 3 $out/faithful_base_n1.err
