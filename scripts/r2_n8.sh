#!/bin/bash
# Round-2 validation on N GPUs of one box (default 8): multi-GPU test tier (fused / weight_avg), flagship bench
# (bucketed, bf16 wire, single launch), K1 trace, fp32 batch-100 and widened+Adam configs, allreduce sweep.
set -u
N=${1:-8}
out=gpurun_out/r2_n$N
mkdir -p $out
if [ "${2:-all}" = all ]; then
timeout 420 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "fused or weight" > $out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/summary.txt
tail -n 12 $out/pytest.log
fi
run() { timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
b="bench.py --gpus $N --steps 300 --warmup 20"
run scripts/k1_trace.py 2>&1 | grep "^{" | tee $out/trace_default.txt
run $b > $out/bench_default.json 2> $out/bench_default.err
run $b --no-e2e --comm-dtype bf16 > $out/bench_bf16.json 2>&1
run $b --no-e2e --no-overlap > $out/bench_single.json 2>&1
if [ "${2:-all}" = all ]; then
run $b --no-e2e --batch 100 --dtype fp32 > $out/bench_fp32_b100.json 2>&1
run $b --no-e2e --batch 100 > $out/bench_bf16_b100.json 2>&1
run bench.py --gpus $N --steps 30 --warmup 5 --model wide --optimizer adam --batch 256 --rows 4096 --no-e2e > $out/bench_wide_adam.json 2>&1
fi
python scripts/benchsum.py $out/bench_*.json | tee -a $out/summary.txt
timeout 200 python -m mpi_tensorflow_b200.parallel.launch -np $N --no-prefix bench/allreduce_sweep.py --max-mb 64 --out $out/sweep.jsonl > $out/sweep.log 2>&1
tail -n 4 $out/sweep.log | cut -c1-400
