"""Join ncu's per-SASS 'Instructions Executed' with nvdisasm -gi line info; print the hottest source lines.
usage: ncu_by_line.py <src.csv from ncu --page source --csv> <nvdisasm -gi output> <kernel-substring>"""
import csv, re, sys, collections
src_csv, sass_txt, kname = sys.argv[1:4]
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]
ci = hdr.index("Instructions Executed"); cs = hdr.index("Source"); cst = hdr.index("# Samples")
body = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name": break  # a second captured launch starts here
    body.append(r)
ncu = [(r[cs].strip(), int(r[ci] or 0), int(r[cst] or 0)) for r in body if len(r) > ci]
# parse nvdisasm: within the function section, track "//## File "...", line N" markers (inlined-at chains too)
lines = open(sass_txt).read().split("\n")
infunc = False; cur = None; sass = []; prev_marker = False
for ln in lines:
    if ln.startswith("//--------------------- .text."):
        infunc = kname in ln
        continue
    if not infunc: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        # keep innermost location plus the outermost "inlined at" search_pca.cu line when present
        inner = (m.group(1).split("/")[-1], int(m.group(2)))
        outer = re.findall(r'inlined at "([^"]+)", line (\d+)', ln)
        if not prev_marker:  # consecutive markers list the inline chain innermost-first: keep the first
            cur = (inner, tuple((o[0].split("/")[-1], int(o[1])) for o in outer))
        prev_marker = True
        continue
    m = re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+(.*?);', ln)
    if m:
        sass.append((m.group(1).strip(), cur)); prev_marker = False
print("ncu rows", len(ncu), "sass rows", len(sass))
agg = collections.Counter(); smp = collections.Counter(); agg_outer = collections.Counter()
n = min(len(ncu), len(sass))
for (s1, cnt, samples), (s2, loc) in zip(ncu[:n], sass[:n]):
    key = loc[0] if loc else ("?", 0)
    agg[key] += cnt; smp[key] += samples
    ok = loc[1][-1] if loc and loc[1] else key
    agg_outer[ok] += cnt
tot = sum(agg.values()); tots = sum(smp.values())
print("total inst", tot, "samples", tots)
print("--- by innermost line"); 
for k, v in agg.most_common(28): print(f"{k[0]}:{k[1]:<5} inst {100*v/tot:5.1f}%  samples {100*smp[k]/max(tots,1):5.1f}%")
print("--- by outermost (kernel) line")
for k, v in agg_outer.most_common(12): print(f"{k[0]}:{k[1]:<5} inst {100*v/tot:5.1f}%")
