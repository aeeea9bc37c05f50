"""Per-launch table from an ncu metrics CSV (`ncu --csv --metrics regex:sm__.*tensor.*,gpu__time_duration.sum,
dram__bytes_read.sum,dram__bytes_write.sum,sm__cycles_elapsed.max`): time, tcgen05 FLOPs as counted by
sm__ops_path_tensor_op_utchmma_*, TFLOP/s, fraction of the sustained bf16 peak, tensor-pipe activity, DRAM bytes.
    python tools/ncu_tensor_table.py peak_tflops file.csv [file.csv ...]"""
import collections
import csv
import sys

peak = float(sys.argv[1])
print("| kernel (launch) | us | ncu FLOP (utchmma) | TFLOP/s | of sustained bf16 peak (%.1f) | pipe_tensor active (of elapsed cycles) | SM clock GHz | DRAM read + write MB |" % peak)
print("|---|---|---|---|---|---|---|---|")
for path in sys.argv[2:]:
    rows = [r for r in csv.reader(open(path)) if len(r) > 8]
    hdr = rows[0]
    ii, ki, mi, vi = hdr.index("ID"), hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
    per = collections.OrderedDict()
    for r in rows[1:]:
        d = per.setdefault(r[ii], {"name": r[ki]})
        try:
            d[r[mi]] = float(r[vi].replace(",", ""))
        except ValueError:
            d[r[mi]] = float("nan")
    for i, d in per.items():
        name = d["name"].replace("void ", "").replace("unnamed>::", "").split("(")[0]
        ns = d.get("gpu__time_duration.sum", float("nan"))
        flop = sum(v for k, v in d.items() if k.startswith("sm__ops_path_tensor_op_utchmma") and v == v)
        cyc = d.get("sm__cycles_elapsed.max", float("nan"))
        act = d.get("sm__pipe_tensor_cycles_active.sum", float("nan"))
        mb = (d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)) / 1e6
        tf = flop / ns / 1e3
        print(f"| `{name}` (#{i}) | {ns / 1e3:.1f} | {flop:.4g} | {tf:.0f} | {tf / peak:.2f} | {act / (4 * 148 * cyc):.2f} | {cyc / ns:.2f} | {mb:.0f} |")
