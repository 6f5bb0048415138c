"""Dump the SASS of one kernel from an ncu report with executed counts and source lines (dev tool).
python tools/ncu_sass.py rep.ncu-rep <kernel-regex> [lib.so] [min_count]"""
import csv, glob, io, os, re, subprocess, sys, tempfile
rep, pat = sys.argv[1], re.compile(sys.argv[2])
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[3] if len(sys.argv) > 3 and sys.argv[3] != "-" else os.path.join(ROOT, "gaussian-transformer_b200", "lib", "libgsrast.so")
minc = int(sys.argv[4]) if len(sys.argv) > 4 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
sass = []
for blk in out.split('"Kernel Name",')[1:]:
    name = blk.split("\n", 1)[0]
    if not pat.search(name):
        continue
    rows = list(csv.reader(io.StringIO(blk.split("\n", 1)[1])))
    hdr = rows[0]
    ia, isrc, iex, isamp, ith = hdr.index("Address"), hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Avg. Predicated-On Threads Executed")
    for r in rows[1:]:
        if len(r) == len(hdr) and r[ia].startswith("0x"):
            sass.append((int(r[ia], 16), r[isrc].strip(), int(r[iex] or 0), int(r[isamp] or 0), r[ith]))
    break
base = sass[0][0]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
lines = {}
for cubin in glob.glob(os.path.join(tmp, "*.cubin")):
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
    inside, cur = False, "?"
    for ln in dis.splitlines():
        if ln.startswith("//---") and ".text." in ln:
            inside = bool(pat.search(subprocess.run(["c++filt", ln.split(".text.")[1].split()[0]], capture_output=True, text=True).stdout))
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
for addr, src, n, samp, th in sass:
    if n >= minc:
        print(f"{addr - base:5x} {n:10d} s{samp:5d} t{th:>5s} {lines.get(addr - base, '?'):22s} {src}")
