"""Summarise .ncu-rep files (read here, no GPU) into profiles/*.md: key raw metrics per kernel
plus the top stall reasons from the source page."""
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    res = []
    for vals in rows[2:]:
        d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
        res.append(d)
    return res


def stalls(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = None
    tot = {}
    for r in rows:
        if "Address" in r[:2] and "Source" in r:
            hdr = r
            continue
        if hdr is None or len(r) < len(hdr):
            continue
        for h, v in zip(hdr, r):
            if h.startswith("stall_") and "Not Issued" not in h:
                try:
                    tot[h] = tot.get(h, 0) + int(v or 0)
                except ValueError:
                    pass
    T = sum(tot.values()) or 1
    return [(k, 100.0 * v / T) for k, v in sorted(tot.items(), key=lambda x: -x[1])[:6]]


def main():
    for rep in sys.argv[1:]:
        print(f"## {rep.split('/')[-1]}\n")
        for d in raw(rep):
            name = d.get("Kernel Name", ("?", ""))[0]
            print(f"### `{name[:90]}`\n")
            print("| metric | value | unit |\n|---|---|---|")
            for w in WANT:
                if w in d:
                    print(f"| {w} | {d[w][0]} | {d[w][1]} |")
            print()
        st = stalls(rep)
        if st:
            print("top warp-stall reasons (all kernels of the report): " + ", ".join(f"{k[6:]} {v:.1f}%" for k, v in st) + "\n")


if __name__ == "__main__":
    main()
