#!/bin/bash
# Backward-kernel tuning sweep (N = 8192 x 2048 x 128256): L2 hints, dX unit order, dX K-splits.
# Pass 1: per-kernel device times (torch.profiler inside flce_microbench.py --kernel-times), no profiler tool attached.
# Pass 2: DRAM bytes per launch of the engine kernels under ncu (cold-cache, serialised: bytes only, not times).
set -u
mkdir -p gpurun_out
MB="python benchmarks/flce_microbench.py --tokens 8192 --masked 0.0 --iters 10 --warmup 3 --no-torch --kernel-times"
MB1="python benchmarks/flce_microbench.py --tokens 8192 --masked 0.0 --iters 1 --warmup 0 --no-torch"
CFGS=("B200SFT_L2_HINTS=0 B200SFT_DX_KS_MAJOR=0"
      "B200SFT_L2_HINTS=0 B200SFT_DX_KS_MAJOR=1 B200SFT_DX_MIN_SPLITS=2"
      "B200SFT_L2_HINTS=1 B200SFT_DX_KS_MAJOR=1 B200SFT_DX_MIN_SPLITS=2"
      "B200SFT_L2_HINTS=2 B200SFT_DX_KS_MAJOR=1 B200SFT_DX_MIN_SPLITS=2"
      "B200SFT_L2_HINTS=1 B200SFT_DX_KS_MAJOR=1 B200SFT_DX_MIN_SPLITS=4")
{
for c in "${CFGS[@]}"; do
  echo "=== $c"
  env $c $MB 2>&1 | grep -E "gemm_tcgen05|fused_bwd_ms" | cut -c1-150
done
i=0
for c in "${CFGS[@]}"; do
  echo "=== ncu $c"
  env $c timeout -k 10 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,sm__cycles_elapsed.avg.per_second,lts__t_sector_hit_rate.pct \
      --clock-control none -k regex:gemm_tcgen05 --csv --log-file gpurun_out/r02_tune_ncu_$i.csv $MB1 > /dev/null 2>&1
  python - gpurun_out/r02_tune_ncu_$i.csv <<'PY'
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]; ix = {n: hdr.index(n) for n in ("ID", "Kernel Name", "Metric Name", "Metric Value")}
per = collections.OrderedDict()
for r in rows[1:]:
    per.setdefault(r[ix["ID"]], {"k": r[ix["Kernel Name"]]})[r[ix["Metric Name"]]] = r[ix["Metric Value"]]
for i, d in per.items():
    f = lambda n: float(d.get(n, "0").replace(",", ""))
    print(f"  {i:>2} rd {f('dram__bytes_read.sum')/1e9:6.3f} GB wr {f('dram__bytes_write.sum')/1e9:6.3f} GB  t {f('gpu__time_duration.sum')/1e6:6.3f} ms  "
          f"clk {f('sm__cycles_elapsed.avg.per_second')/1e9:5.3f} GHz  L2hit {f('lts__t_sector_hit_rate.pct'):5.1f}%  {d['k'][:60]}")
PY
  i=$((i+1))
done
} > gpurun_out/r02_bwd_tune.log 2>&1
cat gpurun_out/r02_bwd_tune.log
