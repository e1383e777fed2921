MB1="python benchmarks/flce_microbench.py --tokens 8192 --masked 0.0 --iters 1 --warmup 0 --no-torch"
for c in "B200SFT_K_PINGPONG=0" "B200SFT_K_PINGPONG=3"; do echo "=== ncu $c"; env $c ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum,sm__cycles_elapsed.avg.per_second,lts__t_sector_hit_rate.pct --clock-control none -k regex:gemm_tcgen05 --csv --log-file gpurun_out/pp.csv $MB1 >/dev/null 2>&1; python - <<PY
import csv, collections
rows=[r for r in csv.reader(open("gpurun_out/pp.csv")) if len(r)>10]
h=rows[0]; ix={n:h.index(n) for n in ("ID","Kernel Name","Metric Name","Metric Value")}
per=collections.OrderedDict()
for r in rows[1:]:
    per.setdefault(r[ix["ID"]],{"k":r[ix["Kernel Name"]][30:62]})[r[ix["Metric Name"]][:22]]=r[ix["Metric Value"]]
for i,d in per.items(): print(i, d)
PY
done
