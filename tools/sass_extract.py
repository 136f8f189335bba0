#!/usr/bin/env python
"""Writes profiles/<name>: per kernel of libffk.so, the tensor-core / tensor-memory / TMA / bulk-copy
SASS instructions (counts and one example line each) — the evidence that the dense path is
tcgen05 + TMEM + TMA code. Runs here (no GPU): cuobjdump -sass on the built library.

    python tools/sass_extract.py [profiles/r02_sass_gemm.txt]
"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
PAT = re.compile(r"\b(UTCHMMA(?:\.2CTA)?|UTCBAR(?:\.2CTA)?(?:\.MULTICAST)?|LDTM(?:\.x\d+)?|STTM(?:\.x\d+)?|"
                 r"UTMALDG(?:\.\dD)?(?:\.2CTA)?(?:\.MULTICAST)?|UBLKCP(?:\.[A-Z.]+)?|UTCATOMSWS(?:\.[A-Z0-9.]+)?)\b")


def main(out):
    lib = ROOT / "flexflow-train_b200" / "libffk.so"
    sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True, check=True).stdout
    demangle = lambda n: subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
    lines = ["# SASS evidence (cuobjdump -sass flexflow-train_b200/libffk.so, sm_100a): tensor-core, tensor-memory, TMA and bulk-copy",
             "# instructions per kernel. UTCHMMA = tcgen05.mma (.2CTA = cta_group::2), LDTM/STTM = tcgen05.ld/st (TMEM),",
             "# UTMALDG = cp.async.bulk.tensor (TMA load), UBLKCP = cp.async.bulk, UTCBAR = tcgen05.commit.",
             "# Regenerate: python tools/sass_extract.py", ""]
    name, counts, example = None, collections.Counter(), {}
    total = collections.Counter()

    def flush():
        if name and counts:
            lines.append(demangle(name)[:160])
            lines.append("    " + ", ".join(f"{k} x{v}" for k, v in sorted(counts.items())))
            for k in sorted(example):
                lines.append("      e.g. " + example[k].strip())

    for ln in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            flush()
            name, counts, example = m.group(1), collections.Counter(), {}
            continue
        m = PAT.search(ln)
        if m and "/*" in ln:
            counts[m.group(1)] += 1
            total[m.group(1)] += 1
            example.setdefault(m.group(1).split(".")[0], ln)
    flush()
    lines += ["", "# totals over the library: " + ", ".join(f"{k} {v}" for k, v in sorted(total.items()))]
    Path(out).write_text("\n".join(lines) + "\n")
    print(f"wrote {out}: {sum(total.values())} instructions in {sum(1 for l in lines if l.startswith('    ') and ' x' in l)} kernels")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else str(ROOT / "profiles" / "r02_sass_gemm.txt"))
