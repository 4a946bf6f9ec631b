"""Join ncu's per-SASS counts with nvdisasm -gi line info and aggregate by REGION of the kernel source
(outermost kernel-file line falls in a named range).
usage: ncu_regions.py <src.csv> <nvdisasm -gi output> <kernel-substring> <file> name:lo-hi [name:lo-hi ...]"""
import csv, re, sys, collections
src_csv, sass_txt, kname, kfile = sys.argv[1:5]
import os
INNER = os.environ.get("INNER", "0") == "1"
regions = []
for a in sys.argv[5:]:
    nm, rg = a.split(":"); lo, hi = rg.split("-"); regions.append((nm, int(lo), int(hi)))
rows = list(csv.reader(open(src_csv)))
# locate the kernel's section
start = None
for i, r in enumerate(rows):
    if r and r[0] == "Kernel Name" and kname_ok(r[1]) if False else (r and r[0] == "Kernel Name"):
        if any(s in r[1] for s in kname.split("|")):
            start = i; break
hdr = rows[start + 1]
ci = hdr.index("Instructions Executed"); cst = hdr.index("# Samples")
stall_cols = [(j, h) for j, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
body = []
for r in rows[start + 2:]:
    if r and r[0] == "Kernel Name": break
    body.append(r)
lines = open(sass_txt).read().split("\n")
infunc = False; cur = None; sass = []; prev_marker = False
mangled = sys.argv[3].split("|")[-1]
for ln in lines:
    if ln.startswith("//--------------------- .text."):
        infunc = mangled in ln; continue
    if not infunc: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        inner = (m.group(1).split("/")[-1], int(m.group(2)))
        outer = re.findall(r'inlined at "([^"]+)", line (\d+)', ln)
        if not prev_marker:
            cur = (inner, tuple((o[0].split("/")[-1], int(o[1])) for o in outer))
        prev_marker = True; continue
    m = re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+(.*?);', ln)
    if m:
        sass.append((m.group(1).strip(), cur)); prev_marker = False
print("ncu rows", len(body), "sass rows", len(sass))
agg = collections.Counter(); smp = collections.Counter(); st = collections.defaultdict(collections.Counter)
for r, (s2, loc) in zip(body, sass):
    cnt = int(r[ci] or 0); samples = int(r[cst] or 0)
    chain = ([loc[0]] + list(loc[1])) if loc else []
    kl = [l for f, l in chain if f == kfile]
    line = (kl[0] if INNER else kl[-1]) if kl else -1
    name = "other"
    for nm, lo, hi in regions:
        if lo <= line <= hi: name = nm; break
    agg[name] += cnt; smp[name] += samples
    for j, h in stall_cols:
        try: st[name][h] += int(r[j] or 0)
        except ValueError: pass
tot = sum(agg.values()); tots = sum(smp.values())
print("total inst", tot, "samples", tots)
for nm, v in agg.most_common():
    top = ", ".join(f"{h[6:]} {100*c/max(smp[nm],1):.0f}%" for h, c in st[nm].most_common(4))
    print(f"{nm:<14} inst {100*v/tot:5.1f}%  samples {100*smp[nm]/max(tots,1):5.1f}%   [{top}]")
