"""Read an `ncu --set full --import-source on` report here (no GPU): headline pipe utilisations and the warp-stall
samples aggregated per SASS opcode.  python tools/ncu_stalls.py report.ncu-rep [top]"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_issued.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.per_cycle_active", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "sm__inst_executed_pipe_tensor",
        "sm__pipe_tensor_subpipe", "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct"]


def page(rep, name):
    return subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 12
    raw = page(rep, "raw")
    rows = list(csv.reader(io.StringIO(raw[raw.index('"ID"'):])))
    head, units = rows[0], rows[1]
    for r in rows[2:]:
        print("==", r[head.index("Kernel Name")][:90], "grid", r[head.index("Grid Size")])
        for h, u, v in zip(head, units, r):
            if any(h.startswith(k) or h == k for k in KEYS) or "issue_stalled" in h and "per_issue_active" in h:
                print(f"   {h:95s} {u:8s} {v}")
    src = page(rep, "source")
    rows = list(csv.reader(io.StringIO(src)))
    hi = next(i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r)
    head = rows[hi]
    col = {h: i for i, h in enumerate(head)}
    stalls = [h for h in head if h.startswith("stall_") and "Not Issued" not in h]
    byop = collections.defaultdict(collections.Counter)
    nexec = collections.Counter()
    for r in rows[hi + 1:]:
        if len(r) < len(head) or r[col["Source"]] == "Source":      # second kernel instance: header repeated
            continue
        s = r[col["Source"]].split()
        if not s:
            continue
        op = (s[1] if s[0].startswith("@") else s[0]).split(".")[0]
        nexec[op] += int(r[col["Instructions Executed"]] or 0)
        for st in stalls:
            byop[op][st] += int(r[col[st]] or 0)
    tot = sum(sum(c.values()) for c in byop.values())
    print(f"-- stall samples by opcode (total {tot}), warp instructions executed {sum(nexec.values())}")
    for op, c in sorted(byop.items(), key=lambda kv: -sum(kv[1].values()))[:top]:
        print(f"   {op:10s} executed {nexec[op]:11d}  samples {sum(c.values()):6d} ({100 * sum(c.values()) / max(tot, 1):4.1f} %)  "
              + ", ".join(f"{k[6:]} {v}" for k, v in c.most_common(4)))


main()
