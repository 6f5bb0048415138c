"""Compact per-kernel metric dump from an ncu --set full report: python tools/ncu_metrics.py rep.ncu-rep [regex]"""
import csv
import io
import re
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "lts__t_bytes.sum", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum",
]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
stall = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
seen = set()
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")]
    if name in seen or (pat and not pat.search(name)):
        continue
    seen.add(name)
    print("----", name[:60])
    for w in WANT:
        if w in hdr:
            print(f"  {w:70s} {r[hdr.index(w)]} {units[hdr.index(w)]}")
    st = sorted(((float(r[hdr.index(h)] or 0), h) for h in stall), reverse=True)[:6]
    print("  top stalls:", ", ".join(f"{h[34:-23]}={v:.2f}" for v, h in st))
