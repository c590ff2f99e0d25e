#!/bin/bash
set -u
out=gpurun_out/r2_n2g
mkdir -p $out
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "train or algebra" > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -n 15 $out/pytest.log
run() { timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
run bench.py --gpus 2 --reference-semantics --impl baseline --steps 200 --warmup 10 > $out/faithful_base_n2.json 2> $out/faithful_base_n2.err
grep "^{" $out/faithful_base_n2.json | cut -c1-400; tail -n 3 $out/faithful_base_n2.err | cut -c1-200
run bench.py --gpus 2 --steps 300 --warmup 20 > $out/bench_default.json 2> $out/bench_default.err
python scripts/benchsum.py $out/bench_*.json
