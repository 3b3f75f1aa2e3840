#!/usr/bin/env python
"""Regenerate profiles/sass/blackwell_mnemonics.md: Blackwell-specific SASS mnemonics per kernel of the built library.
    python tools/sass_evidence.py            (needs cuobjdump; no GPU)"""
import collections
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "distributed-llm-training-and-inference-system_b200" / "lib" / "libllmctl_b200.so"
PAT = re.compile(r"\b(UTC[A-Z0-9]*MMA[.\w]*|UTCBAR[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UBLKCP[.\w]*|UTMAREDG[.\w]*|"
                 r"MULTIMEM[.\w]*|USETMAXREG[.\w]*|FFMA2|FADD2|FMUL2|FMNMX3|SYNCS\.ARRIVE[.\w]*)")


def main():
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
    per = collections.OrderedDict()
    cur = None
    it = iter(names)
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = next(it)
            cur = re.sub(r"\(.*", "", cur).replace("void ", "")
            per.setdefault(cur, collections.Counter())
            continue
        if cur is None:
            continue
        m = PAT.search(line)
        if m:
            key = re.sub(r"\.(x\d+)$", r".\1", m.group(1))
            per[cur][key] += 1
    out = ["# SASS evidence: Blackwell-native instructions per kernel", "",
           "`cuobjdump -sass lib/libllmctl_b200.so`, counted per function by `tools/sass_evidence.py` (UTC*MMA = tcgen05.mma — `tmem[...]` A",
           "operands show up as the same mnemonic, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA tensor load/store, UBLKCP =",
           "cp.async.bulk, UTCBAR = tcgen05.commit, MULTIMEM = NVLS multimem.*, USETMAXREG = setmaxnreg, FFMA2/FADD2/FMUL2 =",
           "packed fp32x2 math, FMNMX3 = 3-input min/max).", "", "| kernel | mnemonics |", "|---|---|"]
    for k in sorted(per):
        c = per[k]
        if not c:
            continue
        out.append(f"| `{k}` | " + ", ".join(f"{n} x{v}" for n, v in sorted(c.items())) + " |")
    p = ROOT / "profiles" / "sass" / "blackwell_mnemonics.md"
    p.write_text("\n".join(out) + "\n")
    print(f"{len(out) - 9} kernels -> {p}")


if __name__ == "__main__":
    main()
