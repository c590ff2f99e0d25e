#!/bin/bash
# Round-2 final multi-GPU measurements on N GPUs of one box: test tier, flagship bench (+ end to end), bf16 wire, the
# fp32 batch-100 and widened+Adam configs, the B-faithful arms, the allreduce sweep.  Results: gpurun_out/r2_final_nN/.
set -u
N=${1:-8}
what=${2:-all}
out=gpurun_out/r2_final_n$N
mkdir -p $out
run() { timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + RANDOM % 400)) "$@"; }
if [ "$what" = all ]; then
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "fused" > $out/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $out/summary.txt
tail -n 6 $out/pytest.log
fi
b="bench.py --gpus $N --steps 300 --warmup 20"
run $b > $out/bench_default.json 2> $out/bench_default.err
if [ "$what" = all ]; then
run $b --no-e2e --comm-dtype bf16 > $out/bench_bf16wire.json 2>&1
run $b --no-e2e --batch 100 --dtype fp32 > $out/bench_fp32_b100.json 2>&1
run bench.py --gpus $N --steps 30 --warmup 5 --model wide --optimizer adam --batch 256 --rows 4096 --no-e2e > $out/bench_wide_adam.json 2>&1
run bench.py --gpus $N --reference-semantics --impl ours --steps 200 --warmup 10 > $out/faithful_ours.json 2> $out/faithful_ours.err
run bench.py --gpus $N --reference-semantics --impl baseline --steps 200 --warmup 10 > $out/faithful_base.json 2> $out/faithful_base.err
fi
python scripts/benchsum.py $out/bench_*.json | tee -a $out/summary.txt
for f in $out/faithful_*.json; do [ -f $f ] && grep "^{" $f | cut -c1-420; done
timeout 200 python -m mpi_tensorflow_b200.parallel.launch -np $N --no-prefix bench/allreduce_sweep.py --max-mb 64 --out $out/sweep.jsonl > $out/sweep.log 2>&1
tail -n 2 $out/sweep.log | cut -c1-300
