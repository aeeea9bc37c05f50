"""Per-kernel totals and shares from ncu launch lists (`--metrics gpu__time_duration.sum --csv`).
    python tools/launch_summary.py TITLE=path.csv [TITLE=path.csv ...] > profiles/rNN_launch_summary.md"""
import collections
import csv
import sys

OURS = ("gemm_tc", "ffn_fused", "attention", "embed_ln", "pool_l2", "topk", "normalize", "attn_", "rerank", "move_rows")


def load(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 8]
    hdr = rows[0]
    ki, mi, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
    return [(r[ki], float(r[vi].replace(",", "")) / 1000) for r in rows[1:] if r[mi] == "gpu__time_duration.sum"]


def short(k):
    k = k.replace("void ", "").replace("lsc::<unnamed>::", "").replace("unnamed>::", "").replace("(lsc::GemmEpilogue)", "")
    d = 0
    for i, ch in enumerate(k):
        if ch == "<":
            d += 1
        elif ch == ">":
            d -= 1
        elif ch == "(" and d == 0:
            return k[:i]
    return k


print("# ncu launch lists (`--metrics gpu__time_duration.sum --clock-control none`; cold cache, serialised: compare SHARES)\n")
for arg in sys.argv[1:]:
    title, path = arg.rsplit("=", 1)
    agg = collections.OrderedDict()
    for k, us in load(path):
        s = short(k)
        if not any(x in s for x in OURS):
            continue
        a = agg.setdefault(s, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(v[1] for v in agg.values())
    print(f"## {title}\n")
    print("| kernel | launches | total us | us per launch | share |\n|---|---|---|---|---|")
    for s, (n, us) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"| `{s}` | {n} | {us:.1f} | {us / n:.1f} | {100 * us / tot:.1f}% |")
    print(f"\ntotal {tot:.1f} us\n")
