#!/bin/bash
# Round-2 multi-GPU measurement pass: bash benchmarks/r02_multi_gpu.sh N [extra]   (run under gpurun --gpus N)
#   headline config 3 (tied-gradient split on = default with the NVLink hook), the same with the split off (N=8 or
#   "extra"), BASELINE config 5 (packed 4096-token rows).  Outputs -> gpurun_out/.
set -u
N=$1; EXTRA=${2:-}
mkdir -p gpurun_out
O=gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 5 --warmup 3"
CPU="--no-cpu-baseline"; { [ "$EXTRA" = "cpu" ] || [ "$EXTRA" = "cpuonly" ]; } && CPU=""
$RUN $CPU > $O/r02_bench_${N}gpu.json 2> $O/r02_bench_${N}gpu.err
$RUN --no-cpu-baseline --workload packed --batch 2 --seq 4096 > $O/r02_bench_packed_${N}gpu.json 2> $O/r02_bench_packed_${N}gpu.err
if [ "$N" = "8" ] || [ "$EXTRA" = "off" ]; then
  $RUN --no-cpu-baseline --tied-exchange off > $O/r02_bench_${N}gpu_tied_off.json 2> $O/r02_bench_${N}gpu_tied_off.err
fi
if [ "$EXTRA" = "cpu" ]; then
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > $O/r02_reference_${N}gpu.json 2> $O/r02_reference_${N}gpu.err
fi
python - <<PY
import json, glob
for f in sorted(glob.glob("$O/r02_bench*${N}gpu*.json") + glob.glob("$O/r02_reference_${N}gpu.json")):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, "NO JSON", e); print(open(f.replace(".json", ".err")).read()[-1500:]); continue
    c = d.get("comm") or {}
    print(f, d.get("value"), d.get("ms_per_step"), "e2e", (d.get("e2e") or {}).get("value"), "exposed", c.get("exposed_ms_per_step"),
          "largest", c.get("largest_bucket"), "tied", json.dumps(c.get("tied_rows", {}).get("max_over_ranks")), json.dumps(c.get("tied_rows", {}).get("min_over_ranks")),
          "valid/rank", (d.get("config") or {}).get("valid_tokens_per_rank"), "cpu", (d.get("cpu_baseline") or {}).get("value"), (d.get("cpu_baseline") or {}).get("cores"))
PY
