"""Executed warp-instructions and stall samples per CUDA source line of one kernel: joins the SASS page of an
ncu report with nvdisasm's line info of the cubin that was profiled (development aid).
usage: python tools/ncu_lines.py REPORT.ncu-rep KERNEL_REGEX CUBIN [top]"""
import csv, re, subprocess, sys, collections

rep, kern, cubin = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 50
sass = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}"], capture_output=True, text=True).stdout
rows = list(csv.reader(sass.splitlines()))
h = next(i for i, r in enumerate(rows) if "Address" in r)
hdr = rows[h]
ia, ii, isamp = hdr.index("Address"), hdr.index("Instructions Executed"), hdr.index("# Samples")
body = []
for r in rows[h + 1:]:
    if len(r) <= ii or not r[ia].startswith("0x"):
        if body:
            break  # second launch of the same kernel
        continue
    body.append((int(r[ia], 16), int(r[ii]), int(r[isamp]), r[hdr.index("Source")]))
base = body[0][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
line_of, cur, infn = {}, None, False
for ln in dis.splitlines():
    if ln.startswith(".text.") or ".section" in ln:
        infn = re.search(kern, ln) is not None
    m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", ln)
    if m and infn:
        line_of[int(m.group(1), 16)] = cur
per = collections.defaultdict(lambda: [0, 0])
tot = samp = 0
for a, n, s, src in body:
    k = line_of.get(a - base, ("?", 0))
    per[k][0] += n
    per[k][1] += s
    tot += n
    samp += s
print(f"total warp-instructions {tot}, samples {samp}")
for k, (n, s) in sorted(per.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{k[0]}:{k[1]:<5} inst {n:>11} {100*n/tot:5.1f}%   samples {s:>7} {100*s/max(samp,1):5.1f}%")
