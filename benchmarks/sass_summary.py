#!/usr/bin/env python
"""Per-kernel counts of the Blackwell-only SASS mnemonics in libb200sft.so: tcgen05.mma (UTCHMMA, .2CTA = cta_group::2),
tcgen05.ld (LDTM), TMA loads / stores (UTMALDG / UTMASTG), tcgen05.commit (UTCBAR), multimem (.MMEM / multicast
load-reduce and stores) -- evidence that the hot kernels are tcgen05 / TMA code, not recompiled mma.sync kernels.

    python benchmarks/sass_summary.py > profiles/r02_sass_summary.txt      (no GPU needed: cuobjdump reads the .so)
"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "llm_fine_tune_distributed_b200", "libb200sft.so")
PATTERNS = collections.OrderedDict([
    ("UTCHMMA", r"\bUTCHMMA"), ("UTCHMMA.2CTA", r"\bUTCHMMA\.2CTA"), ("LDTM", r"\bLDTM"), ("UTMALDG", r"\bUTMALDG"),
    ("UTMASTG", r"\bUTMASTG"), ("UTCBAR", r"\bUTCBAR"), ("multimem", r"\.MMEM|MULTIMEM|\bLDGMC|\bSTGMC|\bREDGMC"),
    ("HMMA/mma.sync", r"\bHMMA\b|\bHMMA\."),
])


def main():
    sass = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True,
                           text=True).stdout.splitlines()
    per, cur, k = collections.OrderedDict(), None, 0
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = names[k]
            k += 1
            per[cur] = collections.Counter()
            continue
        if cur is None or not re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
            continue
        per[cur]["total"] += 1
        for key, pat in PATTERNS.items():
            if re.search(pat, line):
                per[cur][key] += 1
    print(f"# command: cuobjdump -sass {os.path.relpath(SO, ROOT)}  (counted by benchmarks/sass_summary.py)")
    print("# " + " | ".join(["kernel"] + list(PATTERNS) + ["total instructions"]))
    for name in sorted(per):
        c = per[name]
        short = re.sub(r"\(CUtensorMap_st.*", "(...)", name)[:150]
        print(" | ".join([short] + [str(c[key]) for key in PATTERNS] + [str(c["total"])]))
    eng = [n for n in per if "gemm_tcgen05" in n]
    print(f"# {len(eng)} engine instantiations; every one has UTCHMMA > 0: {all(per[n]['UTCHMMA'] > 0 for n in eng)}; "
          f"kernels with HMMA (mma.sync): {sum(1 for n in per if per[n]['HMMA/mma.sync'] > 0)}")


if __name__ == "__main__":
    main()
