"""Condense one `ncu --set full --import-source on` capture (exported with --page details / --page source --csv,
see scratch/ncu_kernel.sh) into a text summary: headline metrics, stall-reason totals, the most-stalled SASS lines."""
import csv
import re
import sys


def main(details, source, out, top=30):
    lines = []
    keep = re.compile(r"Duration|Elapsed Cycles|DRAM Throughput|L2 Cache Throughput|Registers Per Thread|Issue Slots Busy|"
                      r"Executed Ipc Active|Theoretical Occupancy|Dynamic Shared Memory|Block Size|Grid Size|Eligible Warps|No Eligible")
    for ln in open(details):
        if keep.search(ln):
            lines.append(ln.rstrip())
    rows = list(csv.reader(open(source)))
    name = rows[0][1]
    hdr, body = rows[1], rows[2:]
    si, ex = hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
    idx = {n: hdr.index(n) for n in hdr if n.startswith("stall_") and "Not Issued" not in n}
    tot = sum(int(r[si] or 0) for r in body)
    agg = sorted(((sum(int(r[i] or 0) for r in body), n) for n, i in idx.items()), reverse=True)
    with open(out, "w") as f:
        f.write(f"# {name}\n# {len(body)} SASS instructions, {tot} warp-state samples\n")
        f.write("\n".join(lines) + "\n\n# stall reasons (samples)\n")
        for v, n in agg:
            if v:
                f.write(f"{n[6:]:>20} {v:7d} {100.0 * v / max(tot, 1):5.1f} %\n")
        f.write(f"\n# {top} most-stalled instructions: index, samples, times executed, SASS, main reasons\n")
        order = sorted(range(len(body)), key=lambda i: -int(body[i][si] or 0))[:top]
        for i in sorted(order):
            why = {n[6:]: int(body[i][j] or 0) for n, j in idx.items() if int(body[i][j] or 0) > 20}
            f.write(f"{i:5d} {body[i][si]:>6} {body[i][ex]:>9}  {body[i][1][:70]:<70} {why}\n")


if __name__ == "__main__":
    main(*sys.argv[1:4])
