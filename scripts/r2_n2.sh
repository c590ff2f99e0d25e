#!/bin/bash
# Round-2 two-GPU check: multi-GPU test tier, flagship bench bucketed vs single launch, hidden-bucket shape sweep.
set -u
out=gpurun_out/r2_n2
mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/summary.txt
tail -n 25 $out/pytest.log
run() { timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
b="bench.py --gpus 2 --steps 200 --warmup 20 --no-e2e"
run $b > $out/bench_default.json 2> $out/bench_default.err
run $b --no-overlap > $out/bench_single.json 2>&1
MTB_K1_BLOCK_A=32 run $b > $out/bench_b32.json 2>&1
MTB_K1_BLOCK_A=128 run $b > $out/bench_b128.json 2>&1
MTB_COMM_PRIO=-1 run $b > $out/bench_prio_hi.json 2>&1
MTB_K1_INNER_FLAGS=1 MTB_K1_GRID_A=296 run $b > $out/bench_innerflags296.json 2>&1
run $b --comm-algo nvls > $out/bench_nvls.json 2>&1
run $b --comm-dtype bf16 > $out/bench_bf16wire.json 2>&1
run bench.py --gpus 2 --steps 200 --warmup 20 > $out/bench_e2e.json 2>&1
timeout 100 python bench.py --steps 200 --warmup 20 --no-e2e > $out/bench_n1.json 2>&1
python scripts/benchsum.py $out/bench_*.json | tee -a $out/summary.txt
