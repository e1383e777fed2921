#!/bin/bash
# Round-2 one-GPU measurement pass (run under gpurun): config-4 sweep, workload variants of the 1-GPU step, the same
# step on the stock GPU path at growing per-device batch, ncu launch lists and full captures.  Outputs -> gpurun_out/.
set -u
mkdir -p gpurun_out
O=gpurun_out
# (1) BASELINE config 4: masked {0,.25,.5,.75,.9} x {scatter, prefix} x N {8k,16k,32k,64k}, dense and compacted, vs torch
python benchmarks/flce_microbench.py --tokens 8192 16384 32768 65536 --masked 0 0.25 0.5 0.75 0.9 --patterns scatter prefix \
    --modes dense compact --iters 20 --warmup 3 --clocks > $O/r02_flce_microbench_full.jsonl 2> $O/r02_flce_microbench_full.err
# (2) config 2 variants on one GPU: reference-collator padding and completion-only (60 % prompt)
for wl in padded completion; do
  python bench.py --steps 5 --warmup 3 --workload $wl --no-cpu-baseline --no-stock-gpu > $O/r02_bench_1gpu_$wl.json 2> $O/r02_bench_1gpu_$wl.err
done
# (3) what the freed memory buys: both arms at the reference's deployed per-device batch (28) and beyond, until OOM
for b in 28 64 128 256; do
  timeout -k 10 600 python bench.py --steps 2 --warmup 3 --batch $b --no-cpu-baseline --no-stock-gpu --no-e2e > $O/r02_batch_b200_$b.json 2> $O/r02_batch_b200_$b.err || echo "{\"impl\": \"b200\", \"batch\": $b, \"failed\": \"$(grep -o 'OutOfMemoryError' $O/r02_batch_b200_$b.err | head -1)\"}" > $O/r02_batch_b200_$b.json
done
for b in 28 64 80 96; do
  timeout -k 10 600 python bench.py --impl stock --steps 2 --warmup 3 --batch $b --no-cpu-baseline --no-e2e > $O/r02_batch_stock_$b.json 2> $O/r02_batch_stock_$b.err || echo "{\"impl\": \"stock\", \"batch\": $b, \"failed\": \"$(grep -o 'OutOfMemoryError' $O/r02_batch_stock_$b.err | head -1)\"}" > $O/r02_batch_stock_$b.json
done
# (4) ncu: launch list + full capture of the engine kernels (microbench), launch list of this library's kernels in the step
MB="python benchmarks/flce_microbench.py --tokens 8192 --masked 0.0 --iters 2 --warmup 1 --no-torch"
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-clocks --no-stock-gpu"
OURS='regex:gemm_tcgen05|flce_|label_prep|label_frontend|gather_rows|scatter_rows|rmsnorm_|dx_combine|grad_sumsq|grad_norm_finish|adamw_clip|scale_bf16|allreduce_scale|eg_'
$MB > $O/plain_micro.log 2>&1 &&
timeout -k 10 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02_launches_micro.csv $MB > $O/ncu_micro.log 2>&1
$MB > $O/plain_micro2.log 2>&1 &&
timeout -k 10 400 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 2 -c 7 -f -o $O/r02_prof_flce $MB > $O/ncu_full.log 2>&1
$BENCH > $O/plain_bench.log 2>&1 &&
timeout -k 10 420 ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" --csv --log-file $O/r02_launches_bench_ours.csv $BENCH > $O/ncu_bench.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-stock-gpu --kernel-table $O/r02_step_kernels.md > $O/bench_table.log 2>&1
ls -la $O | tail -40
