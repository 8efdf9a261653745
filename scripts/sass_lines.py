"""Histogram of SASS instructions by source file:line for one kernel (nvdisasm -g): where the code size goes."""
import re, subprocess, sys, os, collections, tempfile
so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "python-raytracer_b200", "csrc", "libb200rt.so")
pat = sys.argv[1] if len(sys.argv) > 1 else "k_primaryILb0"
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=d, capture_output=True)
cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
on, cur, hist = False, None, collections.Counter()
for l in txt.splitlines():
    if l.startswith("//-----") and ".text." in l:
        on = pat in l
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', l):
        hist[cur] += 1
tot = sum(hist.values())
print("total", tot)
byfile = collections.Counter()
for (f, ln), c in hist.items():
    byfile[f] += c
print(dict(byfile))
for (f, ln), c in sorted(hist.items(), key=lambda x: -x[1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 40]:
    print(f"{c:6d} {f}:{ln}")
