#!/bin/bash
set -u
out=gpurun_out/r2_n2f
mkdir -p $out
timeout 400 python -m pytest tests/test_gpu_multi.py tests/test_gpu_kernels.py -m gpu -x -q -k "fused or weight" > $out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/summary.txt
tail -n 5 $out/pytest.log
run() { timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
b="bench.py --gpus 2 --steps 200 --warmup 20 --no-e2e"
run scripts/k1_trace.py 2>&1 | grep "^{" | tee $out/trace_default.txt
run $b > $out/bench_default.json 2> $out/bench_default.err
run $b --comm-dtype bf16 > $out/bench_bf16.json 2>&1
timeout 100 python bench.py --steps 200 --warmup 20 --no-e2e > $out/bench_n1.json 2>&1
MTB_OVERLAP_N1=0 timeout 100 python bench.py --steps 200 --warmup 20 --no-e2e > $out/bench_n1_single.json 2>&1
python scripts/benchsum.py $out/bench_*.json | tee -a $out/summary.txt
