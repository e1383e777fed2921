#!/bin/bash
# ncu passes that produced profiles/ (run under gpurun, one GPU).  Each ncu run is preceded by the same
# command without ncu (B200_PROFILING.md).  Outputs land in gpurun_out/; summarise them with
# benchmarks/summarize_ncu.py.  `timeout -k` (TERM, then KILL) so that a stuck run cannot hold the box.
set -u
mkdir -p gpurun_out
MB="python benchmarks/flce_microbench.py --tokens 8192 --masked 0.0 --iters 2 --warmup 1 --no-torch"
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-clocks"
OURS='regex:gemm_tcgen05|flce_|label_prep|label_frontend|gather_rows|scatter_rows|rmsnorm_|dx_combine|grad_sumsq|grad_norm_finish|adamw_clip|scale_bf16|allreduce_scale'

# usage: profile_gpu.sh [micro|step|all]   (default: micro)
WHAT="${1:-micro}"
# (1) microbench: launch list, then full capture of the four engine kernels (fwd, dlogits, dX, dW)
if [ "$WHAT" = "micro" ] || [ "$WHAT" = "all" ]; then
$MB > gpurun_out/plain_micro.log 2>&1 &&
timeout -k 10 200 ncu --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches_micro.csv $MB > gpurun_out/ncu_micro.log 2>&1
$MB > gpurun_out/plain_micro2.log 2>&1 &&
timeout -k 10 300 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 2 -c 4 -f \
    -o gpurun_out/prof_flce $MB > gpurun_out/ncu_full.log 2>&1
fi

# (2) bench.py: launch list restricted to this library's kernels.  Profiling every kernel of the step
#     (tens of thousands of launches) did not finish in 25 minutes; the full per-kernel table of a step
#     comes from `bench.py --kernel-table` (torch.profiler / CUPTI) instead.
if [ "$WHAT" = "step" ] || [ "$WHAT" = "all" ]; then
$BENCH > gpurun_out/plain_bench.log 2>&1 &&
timeout -k 10 420 ncu --metrics gpu__time_duration.sum --clock-control none -k "$OURS" --csv \
    --log-file gpurun_out/launches_bench_ours.csv $BENCH > gpurun_out/ncu_bench.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --kernel-table gpurun_out/step_kernels.md > gpurun_out/bench_table.log 2>&1
fi
ls -la gpurun_out/
