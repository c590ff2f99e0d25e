#!/bin/bash
# Round-1 late additions measured at N GPUs of one box (fp32 config of BASELINE.json, 1 GB allreduce point,
# flagship end-to-end after the loss mailbox):
#   gpurun --gpus N --timeout 600 -- 'bash scripts/final_fp32.sh N'
N=${1:-8}
out=gpurun_out/late_n$N
mkdir -p $out
run() { timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}" 2>$out/err_$1.log | grep '^{' | tail -1; }
run 29521 --steps 300 --warmup 20 --batch 100 --dtype fp32                  > $out/ours_fp32_b100.json
run 29522 --steps 200 --warmup 20 --batch 100 --dtype fp32 --impl baseline  > $out/base_fp32_b100.json
run 29523 --steps 300 --warmup 20                                           > $out/ours_b1024.json
timeout 150 python -m mpi_tensorflow_b200.parallel.launch -np $N --no-prefix bench/allreduce_sweep.py --min-bytes 1073741824 --max-mb 1024 --out $out/sweep_1gb.jsonl > $out/sweep.log 2>&1
for f in $out/*.json; do echo "$f: $(cut -c1-200 $f)"; done
cat $out/sweep_1gb.jsonl | cut -c1-400
tail -2 $out/err_29521.log
