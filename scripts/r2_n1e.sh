#!/bin/bash
# Round-2 evidence run on one GPU: B-faithful arms, compute-sanitizer logs, ncu --set full of the new kernels.
set -u
out=gpurun_out/r2_n1e
mkdir -p $out
timeout 150 python bench.py --reference-semantics --impl ours --steps 200 --warmup 10 2>&1 | grep "^{" | tee $out/faithful_ours_n1.json
timeout 150 python bench.py --reference-semantics --impl baseline --steps 200 --warmup 10 2>&1 | grep "^{" | tee $out/faithful_base_n1.json
# sanitizers on the smallest cases that exercise the new code (memcheck: halo / out-of-bounds boxes of the convolution
# producer and the split-K partials; racecheck: K1's closing warp / last-arriver, split-K last-arriver)
timeout 200 compute-sanitizer --tool memcheck --error-exitcode 99 --log-file $out/sanitizer_memcheck.log \
  python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -p no:cacheprovider -k "conv5x5_implicit and 3-32 or fused_update or conv2_fprop_pool_implicit_gemm and 3" > $out/sanitizer_memcheck.out 2>&1
echo "memcheck rc=$?" | tee -a $out/summary.txt; tail -n 3 $out/sanitizer_memcheck.log
timeout 200 compute-sanitizer --tool racecheck --error-exitcode 99 --log-file $out/sanitizer_racecheck.log \
  python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -p no:cacheprovider -k "fused_update or (gemm_bf16_all_layouts and shape9)" > $out/sanitizer_racecheck.out 2>&1
echo "racecheck rc=$?" | tee -a $out/summary.txt; tail -n 3 $out/sanitizer_racecheck.log
# ncu --set full: the fp32 step's implicit-conv GEMMs and both K1 kernels (only after the plain command exited 0)
cmd="python bench.py --steps 2 --warmup 3 --no-e2e --no-graph --dtype fp32 --batch 100"
timeout 100 $cmd > $out/plain.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k "regex:gemm_bf16_kernel|fused_|splitk" -s 36 -c 12 -o $out/prof_fp32 $cmd > $out/ncu_fp32.log 2>&1
echo "ncu fp32 rc=$?" | tee -a $out/summary.txt
cmd="python bench.py --steps 2 --warmup 3 --no-e2e --no-graph"
timeout 100 $cmd > $out/plain2.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k "regex:fused_" -s 8 -c 4 -o $out/prof_k1 $cmd > $out/ncu_k1.log 2>&1
echo "ncu k1 rc=$?" | tee -a $out/summary.txt
ls -la $out
