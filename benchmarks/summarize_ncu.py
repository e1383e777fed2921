"""Summarise ncu outputs into small tracked files under profiles/.

    python benchmarks/summarize_ncu.py launches <launches.csv> <out.md>
    python benchmarks/summarize_ncu.py full <report.ncu-rep> <out.md>
"""
import csv
import io
import subprocess
import sys
from collections import OrderedDict

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active % (of active cycles)"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe active % (of elapsed)"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2->SM bytes"),
    ("sm__cycles_elapsed.avg.per_second", "SM clock"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
]


def short(name: str) -> str:
    name = name.replace("b200::", "").replace("void ", "")
    return name.split("(")[0][:90]


def launches(path, out):
    rows = [l for l in open(path) if l.startswith('"')]
    r = list(csv.DictReader(io.StringIO("".join(rows))))
    agg = OrderedDict()
    total = 0.0
    for x in r:
        v = float(x["Metric Value"].replace(",", ""))
        unit = x["Metric Unit"]
        ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)
        k = short(x["Kernel Name"])
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += ns
        total += ns
    with open(out, "w") as f:
        f.write(f"# ncu launch list summary ({path})\n\n`ncu --metrics gpu__time_duration.sum --clock-control none` -- "
                "cold-cache, serialised launches: compare shares, not absolutes.\n\n")
        f.write("| kernel | launches | total ms | avg us | share |\n|---|---:|---:|---:|---:|\n")
        for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"| `{k}` | {n} | {ns / 1e6:.3f} | {ns / n / 1e3:.1f} | {100 * ns / total:.1f}% |\n")
        f.write(f"\ntotal {total / 1e6:.3f} ms over {len(r)} launches\n")


def full(path, out):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    r = list(csv.reader(io.StringIO(raw)))
    hdr, units = r[0], r[1]
    col = {h: i for i, h in enumerate(hdr)}
    with open(out, "w") as f:
        f.write(f"# ncu --set full summary ({path})\n\n`ncu --set full --clock-control none --import-source on`, one row "
                "per captured launch (values per launch).\n\n")
        for row in r[2:]:
            f.write(f"## `{short(row[col['Kernel Name']])}`\n\n| metric | value |\n|---|---|\n")
            for key, label in KEYS:
                if key in col:
                    f.write(f"| {label} (`{key}`) | {row[col[key]]} {units[col[key]]} |\n")
            f.write("\n")


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2], sys.argv[3])
