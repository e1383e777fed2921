#!/bin/bash
# ncu passes for profiles/ (run under gpurun, one GPU).  Each ncu run is preceded by the same
# command without ncu (B200_PROFILING.md).  Outputs land in gpurun_out/.
set -u
mkdir -p gpurun_out
MB="python benchmarks/flce_microbench.py --tokens 8192 --masked 0.0 --iters 2 --warmup 1 --no-torch"
BENCH="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"

# (1) microbench: launch list + full capture of the engine kernels
$MB > gpurun_out/plain_micro.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_micro.csv $MB > gpurun_out/ncu_micro.log 2>&1
$MB > gpurun_out/plain_micro2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05_kernel -s 4 -c 4 -f -o gpurun_out/prof_flce $MB > gpurun_out/ncu_full.log 2>&1

# (2) whole step: launch list of bench.py (share of the step taken by each kernel)
if [ "${1:-}" = "step" ]; then
$BENCH > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_bench.csv $BENCH > gpurun_out/ncu_bench.log 2>&1
fi
ls -la gpurun_out/
