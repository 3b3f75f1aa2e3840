#!/usr/bin/env python
"""Regenerate profiles/sass/: (1) blackwell_mnemonics.md — Blackwell-specific SASS mnemonics per kernel of the built
library, (2) one full SASS listing per fused compute+collective / peer-memory kernel (K1, K2, K7, K11, collectives).
    python tools/sass_evidence.py            (needs cuobjdump; no GPU)

NVLS shows up as LDGMC (multimem.ld_reduce: `LDGMC.E.ADD.F32x...`/`.HPADD.BF16x8...`), REDG...MC / STG on a multicast
address (multimem.red / multimem.st compile to `STG.E.128.STRONG.SYS` — the multicast mapping is in the address, not in
the opcode); peer signalling as LD/ST ... .STRONG.SYS and MEMBAR/FENCE at system scope."""
import collections
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "distributed-llm-training-and-inference-system_b200" / "lib" / "libllmctl_b200.so"
PAT = re.compile(r"\b(UTC[A-Z0-9]*MMA[.\w]*|UTCBAR[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UBLKCP[.\w]*|UTMAREDG[.\w]*|"
                 r"MULTIMEM[.\w]*|LDGMC[.\w]*|REDG[.\w]*|ATOMG[.\w]*SYS[.\w]*|STG\.E[.\w]*STRONG\.SYS|LDG\.E[.\w]*STRONG\.SYS|LD\.E[.\w]*STRONG\.SYS|"
                 r"ST\.E[.\w]*STRONG\.SYS|MEMBAR[.\w]*SYS|FENCE[.\w]*|USETMAXREG[.\w]*|FFMA2|FADD2|FMUL2|FMNMX3|SYNCS\.ARRIVE[.\w]*)")

#: kernels whose complete SASS is committed (first template instance of each)
LISTINGS = ["ag_push_kernel", "gemm2_tcgen05_kernel<false, false, __nv_bfloat16>", "gemm_tcgen05_kernel<256, false, false, __nv_bfloat16>",
            "gemm_rs_kernel<false>", "rs_reduce_rmsnorm_kernel<16>", "rs_sumsq_kernel<true>", "rs_sumsq_kernel<false>",
            "adam_allgather_kernel", "ar_residual_rmsnorm_kernel<2>", "allreduce_multimem_kernel", "allgather_pull_kernel",
            "reduce_scatter_pull_kernel", "allreduce_oneshot_kernel", "signal_wait_kernel"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
    per = collections.OrderedDict()
    text = collections.OrderedDict()
    cur = None
    it = iter(names)
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = next(it)
            cur = re.sub(r"\(.*", "", cur).replace("void ", "")
            per.setdefault(cur, collections.Counter())
            text.setdefault(cur, [])
            continue
        if cur is None:
            continue
        text[cur].append(line)
        m = PAT.search(line)
        if m:
            key = re.sub(r"\.(x\d+)$", r".\1", m.group(1))
            per[cur][key] += 1
    out = ["# SASS evidence: Blackwell-native instructions per kernel", "",
           "`cuobjdump -sass lib/libllmctl_b200.so`, counted per function by `tools/sass_evidence.py` (UTC*MMA = tcgen05.mma — `tmem[...]` A",
           "operands show up as the same mnemonic, LDTM/STTM = tcgen05.ld/st, UTMALDG/UTMASTG = TMA tensor load/store, UBLKCP =",
           "cp.async.bulk, UTCBAR = tcgen05.commit, LDGMC = NVLS multimem.ld_reduce, STG...STRONG.SYS = multimem.st / system-scope",
           "release stores, LD/LDG...STRONG.SYS = acquire loads of peer flags, MEMBAR/FENCE = system fences, USETMAXREG = setmaxnreg,",
           "FFMA2/FADD2/FMUL2 = packed fp32x2 math, FMNMX3 = 3-input min/max).  Full listings of the fused compute+collective",
           "kernels are the `*.sass` files next to this table.", "", "| kernel | mnemonics |", "|---|---|"]
    for k in sorted(per):
        c = per[k]
        if not c:
            continue
        out.append(f"| `{k}` | " + ", ".join(f"{n} x{v}" for n, v in sorted(c.items())) + " |")
    p = ROOT / "profiles" / "sass" / "blackwell_mnemonics.md"
    p.write_text("\n".join(out) + "\n")
    print(f"{len(out) - 9} kernels -> {p}")
    for want in LISTINGS:
        hit = [k for k in text if k.replace("llmctl::", "") == want]
        if not hit:
            print(f"  (no function named {want})")
            continue
        fn = re.sub(r"[^A-Za-z0-9_]+", "_", want).strip("_") + ".sass"
        body = [ln for ln in text[hit[0]] if re.search(r"/\*[0-9a-f]{4}\*/", ln)]
        # keep the instruction column only (address + opcode + operands): the encoding words double the size for nothing
        body = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", ln).rstrip() for ln in body]
        (ROOT / "profiles" / "sass" / fn).write_text(f"// {hit[0]}  (cuobjdump -sass, sm_100a)\n" + "\n".join(body) + "\n")
        print(f"  {fn}: {len(body)} instructions")


if __name__ == "__main__":
    main()
