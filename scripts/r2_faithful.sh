#!/bin/bash
# B-faithful arms (bench.py --reference-semantics) on N GPUs and on one: ours vs the literal library restatement.
set -u
N=${1:-2}
out=gpurun_out/r2_faithful
mkdir -p $out
run() { timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
timeout 150 python bench.py --reference-semantics --impl baseline --steps 200 --warmup 10 > $out/base_n1.json 2> $out/base_n1.err
timeout 150 python bench.py --reference-semantics --impl ours --steps 200 --warmup 10 > $out/ours_n1.json 2> $out/ours_n1.err
run bench.py --gpus $N --reference-semantics --impl baseline --steps 200 --warmup 10 > $out/base_n$N.json 2> $out/base_n$N.err
run bench.py --gpus $N --reference-semantics --impl ours --steps 200 --warmup 10 > $out/ours_n$N.json 2> $out/ours_n$N.err
for f in $out/*.json; do echo "$f: $(grep '^{' $f | cut -c1-400)"; done
tail -n 6 $out/base_n1.err
