"""Stall samples / executed instructions per CUDA source line of one kernel: joins the SASS rows of
`ncu -i REP --page source --csv` with the line table of `nvdisasm -g -c` on the cubin extracted from the
library that produced the report (cuobjdump -xelf all lib/libvistaf.so).
  [NCU_FILTER='--kernel-name-base demangled -k regex:IllumStore'] python tools/ncu_lines.py REP.ncu-rep CUBIN MANGLED_KERNEL_SUBSTRING [top]"""
import collections, csv, io, re, subprocess, sys
rep, cubin, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
sass = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
seq, cur, infn = [], None, False
for l in sass.split("\n"):
    if l.startswith(".text.") and kern in l and l.rstrip().endswith(":"):
        infn = True
        continue
    if infn and l.startswith("//---"):
        break
    if not infn:
        continue
    m = re.search(r'//## File ".*?/([^/"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+.*?;", l):
        seq.append(cur)
import os
extra = os.environ.get("NCU_FILTER", "").split()
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + extra, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, data = rows[1], rows[2:]
si, ie = hdr.index("# Samples"), hdr.index("Instructions Executed")
assert len(seq) == len(data), (len(seq), len(data))
by = collections.defaultdict(lambda: [0, 0])
for ln, r in zip(seq, data):
    by[ln][0] += int(r[si]); by[ln][1] += int(r[ie])
ts, ti = sum(v[0] for v in by.values()), sum(v[1] for v in by.values())
print(f"samples {ts}  warp instructions {ti}")
for ln, v in sorted(by.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{ln[0] if ln else '?'}:{ln[1] if ln else 0:<5d} {100*v[0]/ts:5.1f}% samples {100*v[1]/ti:5.1f}% instr")
