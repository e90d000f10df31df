"""Turn gpurun_out/*.ncu-rep (+ the bench launch list) into small text summaries under profiles/."""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.avg",
        "smsp__cycles_active.avg", "smsp__inst_executed.avg.per_cycle_active"]


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    return hdr, units, rows[2:]


def stalls(rep, top=int(__import__("os").environ.get("TOP", "14"))):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h = rows[1]
    data = []
    for r in rows[2:]:  # first kernel of the report only (the header repeats per kernel)
        if len(r) != len(h):
            continue
        if r == h:
            break
        data.append(r)
    iS, iSrc, iE = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
    st = [i for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    tot = sum(int(r[iS]) for r in data)
    lines = [f"warp-state samples: {tot}"]
    agg = sorted(((h[j], sum(int(r[j]) for r in data)) for j in st), key=lambda kv: -kv[1])[:8]
    lines.append("stall reasons (all samples): " + ", ".join(f"{k}={v}" for k, v in agg))
    lines.append("hottest SASS instructions (samples, executed, instruction, top stalls):")
    for i in sorted(sorted(range(len(data)), key=lambda i: -int(data[i][iS]))[:top]):
        r = data[i]
        s = sorted(((h[j], int(r[j])) for j in st if int(r[j]) > 0), key=lambda kv: -kv[1])[:2]
        lines.append(f"  {r[iS]:>5} {r[iE]:>8}  {r[iSrc].strip()[:70]:<70} {s}")
    return "\n".join(lines)


def main():
    rep, title, dst = sys.argv[1], sys.argv[2], sys.argv[3]
    hdr, units, rows = raw(rep)
    with open(dst, "w") as f:
        f.write(f"# {title}\n# source: {rep} (ncu --set full --clock-control none --import-source on)\n")
        for r in rows:
            f.write(f"\nkernel: {r[hdr.index('Kernel Name')]}  grid {r[hdr.index('Grid Size')]} block {r[hdr.index('Block Size')]}\n")
            for k in KEYS:
                if k in hdr:
                    f.write(f"  {k:<72} {r[hdr.index(k)]} {units[hdr.index(k)]}\n")
        f.write("\n" + stalls(rep) + "\n")
    print("wrote", dst)


if __name__ == "__main__":
    main()
