#!/bin/bash
set -u
N=${1:-8}
out=gpurun_out/r2_n${N}b
mkdir -p $out
run() { timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
b="bench.py --gpus $N --steps 300 --warmup 20"
run scripts/k1_trace.py 2>&1 | grep "^{" | tee $out/trace_default.txt
run $b > $out/bench_default.json 2> $out/bench_default.err
run $b --no-e2e --comm-dtype bf16 > $out/bench_bf16.json 2>&1
python scripts/benchsum.py $out/bench_*.json | tee -a $out/summary.txt
if [ "${2:-}" = sweep ]; then
timeout 200 python -m mpi_tensorflow_b200.parallel.launch -np $N --no-prefix bench/allreduce_sweep.py --max-mb 64 --out $out/sweep.jsonl > $out/sweep.log 2>&1
tail -n 3 $out/sweep.log | cut -c1-300
fi
