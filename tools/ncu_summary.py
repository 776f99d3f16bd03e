"""Turn `ncu --set full` reports into the small JSON summary bench.py reads (profiles/r01_ncu_full_summary.json).
Run here (no GPU needed):  python tools/ncu_summary.py out.json "label" rep1.ncu-rep [rep2.ncu-rep ...]"""
import csv
import io
import json
import subprocess
import sys

WANT = {  # summary key -> substrings that identify the metric column (first match wins)
    "time_us": ["gpu__time_duration.sum"],
    "dram_read_MB": ["dram__bytes_read.sum"],
    "dram_write_MB": ["dram__bytes_write.sum"],
    "tensor_pipe_pct": ["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
                        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"],
    "xu_pipe_pct": ["sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"],
    "issue_active_pct": ["sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__issue_active.avg.pct"],
    "l2_hit_pct": ["lts__t_sector_hit_rate.pct"],
    "regs": ["launch__registers_per_thread"],
    "warps_active_pct": ["sm__warps_active.avg.pct_of_peak_sustained_active"],
    "smem_wavefront_pct": ["l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed",
                           "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"],
}
SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}


def rows_of(rep):
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    txt = txt[txt.index('"ID"'):]
    rd = list(csv.reader(io.StringIO(txt)))
    return rd[0], rd[1], rd[2:]


def main():
    out_path, label, reps = sys.argv[1], sys.argv[2], sys.argv[3:]
    out = []
    for rep in reps:
        head, units, rows = rows_of(rep)
        col = {h: i for i, h in enumerate(head)}
        for r in rows:
            e = dict(kernel=r[col["Kernel Name"]][:70], grid=r[col["Grid Size"]], source=f"{label} ({rep.split('/')[-1]})")
            for key, pats in WANT.items():
                idx = next((i for p in pats for i, h in enumerate(head) if h == p), None)
                if idx is None:
                    idx = next((i for p in pats for i, h in enumerate(head) if h.endswith(p)), None)
                if idx is None or r[idx] in ("", "n/a"):
                    e[key] = None
                    continue
                v = float(r[idx].replace(",", ""))
                e[key] = round(v * SCALE.get(units[idx], 1.0), 5)
            out.append(e)
    with open(out_path, "w") as f:
        json.dump(out, f, indent=0)
    for e in out:
        print(f"{e['kernel'][:48]:48s} {e['time_us']:9.1f} us  dram {e['dram_read_MB']}+{e['dram_write_MB']} MB  "
              f"tensor {e['tensor_pipe_pct']}  xu {e['xu_pipe_pct']}  regs {e['regs']}")


main()
