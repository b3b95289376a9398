#!/usr/bin/env python
"""Regenerate the SASS evidence under profiles/ from the built library (no GPU needed):

  profiles/sass_summary.txt            per kernel: instruction count, registers and the Blackwell / NVLink opcodes
  profiles/sass/all_kernels.sass.xz    the complete `cuobjdump -sass` listing of every kernel (xz, ~20x smaller)
  profiles/sass/<kernel>.txt           plain listings of the kernels named in KEEP

    python scripts/sass_evidence.py
"""
from __future__ import annotations

import collections
import lzma
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "distributed_training_b200" / "_native" / "libdtb200.so"
OUT = ROOT / "profiles"
# opcodes worth counting (prefix match on the mnemonic)
WATCH = ["UTCHMMA", "UTMALDG", "UTMASTG", "UBLKCP", "UBLKRED", "LDTM", "STTM", "UTCBAR", "UTCATOMSWS", "SYNCS.ARRIVE",
         "SYNCS.PHASECHK", "ELECT", "MUFU.EX2", "LDGMC", "REDG", "STAS", "UCGABAR", "LDG.E.128", "STG.E.128", "STS.128",
         "LDG.E.ENL2.256", "STG.E.ENL2.256", "SHFL", "ATOMG", "MEMBAR", "CCTL"]
KEEP = ["allreduce_nvls_kernel", "broadcast_nvls_kernel", "allgather_nvls_kernel", "bwd_ws_kernel", "bwd_fused_kernel",
        "gather_rows_kernel", "ce_stream_kernel", "fwd_lazy_kernel"]


def main() -> None:
    sass = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    res = subprocess.run(["cuobjdump", "-res-usage", str(LIB)], capture_output=True, text=True, check=True).stdout
    regs = {}
    cur = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
        m = re.search(r"REG:(\d+).*SHARED:(\d+)", line)
        if m and cur:
            regs[cur] = (int(m.group(1)), int(m.group(2)))
    (OUT / "sass").mkdir(parents=True, exist_ok=True)
    with lzma.open(OUT / "sass" / "all_kernels.sass.xz", "wt", preset=6) as fh:
        fh.write(sass)
    blocks = re.split(r"(?=\tFunction : )", sass)
    rows = []
    demangle = subprocess.run(["c++filt"], input="\n".join(regs), capture_output=True, text=True).stdout.splitlines()
    pretty = dict(zip(regs, demangle))
    kept = set()
    for blk in blocks:
        m = re.match(r"\tFunction : (\S+)", blk)
        if not m:
            continue
        name = m.group(1)
        ops = re.findall(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]+)", blk)
        count = collections.Counter()
        for op in ops:
            for w in WATCH:
                if op.startswith(w):
                    count[w] += 1
                    break
        r, sh = regs.get(name, (0, 0))
        short = re.sub(r"\(.*", "", pretty.get(name, name))
        rows.append((short, len(ops), r, sh, count))
        for k in KEEP:
            if k in name and k not in kept:
                kept.add(k)
                (OUT / "sass" / f"{k}.txt").write_text(blk)
    rows.sort()
    with open(OUT / "sass_summary.txt", "w") as fh:
        fh.write("# SASS evidence per kernel in libdtb200.so (cuobjdump -sass / -res-usage, sm_100a); regenerate with\n"
                 "# scripts/sass_evidence.py.  Full listings: profiles/sass/all_kernels.sass.xz (xz -dc to read).\n"
                 "# UTCHMMA = tcgen05.mma, UTMALDG = TMA load, UTMASTG = TMA store, UBLKRED = cp.reduce.async.bulk,\n"
                 "# LDTM = tcgen05.ld, UTCBAR = tcgen05.commit, UTCATOMSWS = TMEM alloc/dealloc, SYNCS.* = mbarrier,\n"
                 "# LDGMC = multimem.ld_reduce (NVLS; multimem.st compiles to a plain STG on the multicast address),\n"
                 "# STAS = st.async, UCGABAR = cluster barrier\n\n")
        for short, n, r, sh, count in rows:
            fh.write(f"{short}   [{n} instr, {r} regs, {sh} B static smem]\n")
            if count:
                fh.write("    " + ", ".join(f"{k}x{v}" for k, v in count.items()) + "\n")
    print(f"{len(rows)} kernels; kept listings: {sorted(kept)}")


if __name__ == "__main__":
    main()
