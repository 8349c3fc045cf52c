#!/usr/bin/env python
"""tools/ncu_summary.py <report.ncu-rep> [out.md] — text summary of one ncu --set full capture:
key raw metrics (duration, DRAM bytes, pipe utilisation, registers) and the per-instruction stall
hot spots from the source page.  The summaries are what gets committed under profiles/."""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second",
        "sm__cycles_elapsed.avg", "sm__cycles_elapsed.avg.per_second", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed"]


def run(args):
    return subprocess.run(["ncu", "-i", *args], capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    out = open(sys.argv[2], "w") if len(sys.argv) > 2 else sys.stdout
    rows = list(csv.reader(io.StringIO(run([rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        out.write(f"## {name[:110]}\n\n| metric | unit | value |\n|---|---|---|\n")
        for h, u, v in zip(hdr, units, r):
            if h in KEYS or "imma" in h and "pct" in h:
                out.write(f"| {h} | {u} | {v} |\n")
        out.write("\n")
    src = list(csv.reader(io.StringIO(run([rep, "--page", "source", "--csv"]))))
    # the source page of the first kernel follows a "Kernel Name" line and a header line
    try:
        hi = next(i for i, r in enumerate(src) if r and r[0] == "Address")
    except StopIteration:
        return
    hdr = src[hi]
    ix = {h: i for i, h in enumerate(hdr)}
    # several kernels in one report repeat the header: keep the kernel with the most samples
    groups, cur = [], []
    for r in src[hi + 1:]:
        if len(r) == len(hdr) and r[ix["# Samples"]].strip().isdigit():
            cur.append(r)
        elif r and r[0] == "Address":
            groups.append(cur); cur = []
    groups.append(cur)
    data = max(groups, key=len)        # the longest kernel = the knn kernel, not its helpers
    stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[ix["# Samples"]] or 0) for r in data) or 1
    agg = sorted(((sum(int(r[ix[s]] or 0) for r in data), s) for s in stalls), reverse=True)
    out.write("### warp stall samples (all instructions)\n\n")
    for v, s in agg[:8]:
        out.write(f"- {s}: {100.0 * v / tot:.1f}%\n")
    out.write("\n### hottest instructions\n\n| samples % | executed | SASS | top stalls |\n|---|---|---|---|\n")
    for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]] or 0))[:25]:
        st = sorted(((int(r[ix[s]] or 0), s) for s in stalls), reverse=True)[:2]
        out.write(f"| {100.0 * int(r[ix['# Samples']] or 0) / tot:.2f} | {r[ix['Instructions Executed']]} | "
                  f"`{r[ix['Source']][:80]}` | {', '.join(f'{s}={v}' for v, s in st)} |\n")


if __name__ == "__main__":
    main()
