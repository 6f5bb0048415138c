"""Per-source-line instruction / stall-sample counts for one kernel (dev tool).

python tools/ncu_lines.py rep.ncu-rep <kernel-regex> [lib.so] [top]
Joins the SASS page of an `ncu --set full --import-source on` report with `nvdisasm -g` line info of the same
library build (instruction order is identical)."""
import csv
import glob
import io
import os
import re
import subprocess
import sys
import tempfile

rep, pat = sys.argv[1], re.compile(sys.argv[2])
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[3] if len(sys.argv) > 3 and sys.argv[3] != "-" else os.path.join(ROOT, "gaussian-transformer_b200", "lib", "libgsrast.so")
top = int(sys.argv[4]) if len(sys.argv) > 4 else 45

out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
sass = []
for blk in out.split('"Kernel Name",')[1:]:
    name = blk.split("\n", 1)[0]
    if not pat.search(name):
        continue
    rows = list(csv.reader(io.StringIO(blk.split("\n", 1)[1])))
    hdr = rows[0]
    ia, isrc, iex, isamp = (hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"),
                            hdr.index("# Samples"))
    for r in rows[1:]:
        if len(r) == len(hdr) and r[ia].startswith("0x"):
            sass.append((int(r[ia], 16), r[isrc].strip(), int(r[iex] or 0), int(r[isamp] or 0)))
    print("====", name[:90])
    break
if not sass:
    sys.exit("kernel not found in report")
base = sass[0][0]

tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
lines = {}
for cubin in glob.glob(os.path.join(tmp, "*.cubin")):
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    inside, cur = False, "?"
    for ln in dis.splitlines():
        if ln.startswith("//---") and ".text." in ln:
            inside = bool(pat.search(subprocess.run(["c++filt", ln.split(".text.")[1].split()[0]],
                                                    capture_output=True, text=True).stdout))
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = f"{os.path.basename(m.group(1))}:{m.group(2)}"
            continue
        m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*?);", ln)
        if m:
            lines[int(m.group(1), 16)] = cur
agg, total, tsamp = {}, 0, 0
for addr, src, n, samp in sass:
    key = lines.get(addr - base, "?")
    a = agg.setdefault(key, [0, 0, 0])
    a[0] += n
    a[1] += samp
    a[2] += 1
    total += n
    tsamp += samp
srccache = {}


def text(key):
    f, _, l = key.partition(":")
    for d in ("gaussian-transformer_b200/csrc",):
        p = os.path.join(ROOT, d, f)
        if os.path.exists(p):
            if p not in srccache:
                srccache[p] = open(p).read().splitlines()
            try:
                return srccache[p][int(l) - 1].strip()
            except Exception:
                return ""
    return ""


print(f"total warp-instructions {total}, samples {tsamp}")
for key, (n, samp, cnt) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{n:12d} {100.0 * n / max(total, 1):5.1f}%  samp {100.0 * samp / max(tsamp, 1):5.1f}%  sass {cnt:3d}  "
          f"{key:24s} {text(key)[:90]}")
