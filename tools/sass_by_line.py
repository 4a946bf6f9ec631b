"""SASS instructions per source line of one kernel (nvdisasm -g), biggest first.
usage: sass_by_line.py <file.o> <kernel-name-substring> [top]"""
import collections, glob, os, re, subprocess, sys, tempfile
obj, key = os.path.abspath(sys.argv[1]), sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd=d, capture_output=True)
    dis = subprocess.run(["nvdisasm", "-g", glob.glob(d + "/*.cubin")[0]], capture_output=True, text=True).stdout
on, cur, cnt = False, "?", collections.Counter()
for ln in dis.splitlines():
    if ln.startswith(".text."):
        on = key in ln
    elif on:
        m = re.search(r'//## File "([^"]*)", line (\d+)', ln)
        if m:
            cur = os.path.basename(m.group(1)) + ":" + m.group(2)
        elif re.match(r"\s+/\*[0-9a-f]+\*/\s", ln):
            cnt[cur] += 1
print(sum(cnt.values()), "TOTAL")
for k, v in cnt.most_common(top):
    print(v, k)
