#!/bin/bash
# All multi-GPU measurements that go into BASELINE.md / profiles, for N GPUs of one box:
#   gpurun --gpus N --timeout 900 -- 'bash scripts/final_multi.sh N'
N=${1:-8}
out=gpurun_out/final_n$N
mkdir -p $out
run() { timeout 250 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}" 2>&1 | grep '^{' | tail -1; }
run 29511 --steps 300 --warmup 20                                      > $out/ours_b1024.json
run 29512 --steps 300 --warmup 20 --batch 100 --no-e2e                 > $out/ours_b100.json
run 29513 --steps 200 --warmup 20 --batch 100 --no-e2e --impl baseline > $out/base_b100.json
run 29514 --steps 30 --warmup 5 --model wide --optimizer adam --batch 256 --rows 4096 --no-e2e > $out/ours_wide_adam.json
run 29515 --steps 30 --warmup 5 --model wide --optimizer adam --batch 256 --rows 4096 --no-e2e --impl baseline > $out/base_wide_adam.json
timeout 250 python -m mpi_tensorflow_b200.parallel.launch -np $N --no-prefix bench/allreduce_sweep.py --max-mb 256 --out $out/sweep.jsonl > $out/sweep.log 2>&1
for f in $out/*.json; do echo "$f: $(cut -c1-260 $f)"; done
tail -3 $out/sweep.jsonl | cut -c1-300
