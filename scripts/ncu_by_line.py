"""Joins an `ncu --page source --csv` SASS profile with `nvdisasm -g` line info and aggregates by source line.
usage: ncu_by_line.py <report.ncu-rep> <kernel substring> <sass file from nvdisasm -g -c> <mangled-name substring> [launch index]"""
import csv, io, re, subprocess, sys, collections
rep, kname, sass, mangled = sys.argv[1:5]
which = int(sys.argv[5]) if len(sys.argv) > 5 else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
# split per kernel
blocks, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "rows": []}; blocks.append(cur)
    elif cur is not None:
        cur["rows"].append(row)
blocks = [b for b in blocks if kname in b["name"]]
b = blocks[which]
hdr = b["rows"][0]; rows = b["rows"][1:]
ci = {h: i for i, h in enumerate(hdr)}
# line info: sequence of (line) per instruction in function order
lines, in_fn, curline = [], False, None
for l in open(sass):
    if l.startswith(".text.") and l.rstrip().endswith(":"):
        in_fn = mangled in l
        continue
    if not in_fn: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        curline = (m.group(1).split("/")[-1], int(m.group(2))); continue
    if re.search(r"/\*[0-9a-f]{4,}\*/", l) and not l.strip().startswith(".") :
        lines.append(curline)
print(f"{b['name'][:80]}: profile rows {len(rows)}, disasm instructions {len(lines)}")
agg = collections.defaultdict(lambda: [0, 0, 0])
for i, r in enumerate(rows):
    ln = lines[i] if i < len(lines) else ("?", 0)
    a = agg[ln]
    a[0] += int(r[ci["Instructions Executed"]]); a[1] += int(r[ci["Thread Instructions Executed"]]); a[2] += int(r[ci["# Samples"]])
tot = sum(a[0] for a in agg.values()); tots = sum(a[2] for a in agg.values())
print(f"total warp insts {tot}, samples {tots}")
srcs = {}
for (f, n), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:45]:
    try:
        if f not in srcs: srcs[f] = open("/root/repo/ignis-masterthesis_b200/csrc/" + f).read().splitlines()
        text = srcs[f][n - 1].strip()[:90]
    except Exception: text = ""
    print(f"{100*a[0]/tot:5.1f}% inst {100*a[2]/max(tots,1):5.1f}% stall  thr/inst {a[1]/max(a[0],1):5.1f}  {f}:{n}  {text}")
