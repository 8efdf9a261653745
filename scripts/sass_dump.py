"""Static SASS of one kernel restricted to instructions whose inline stack passes through rt_core.cuh lines [lo, hi].
usage: sass_dump.py kernel_pattern lo hi [lib.so]"""
import os, re, subprocess, sys, tempfile
pat, lo, hi = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
so = sys.argv[4] if len(sys.argv) > 4 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "python-raytracer_b200", "csrc", "libb200rt.so")
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=d, capture_output=True)
cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
on, stack, pending, n = False, [], [], 0
for l in txt.splitlines():
    if l.startswith("//-----") and ".text." in l:
        on = pat in l; continue
    if not on: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
    if m: pending.append(m); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', l) or re.match(r'\s*\.L_x_\d+:', l):
        if l.strip().startswith(".L_x"):
            print(l.strip()); continue
        if pending:
            st = [(os.path.basename(pending[0].group(1)), int(pending[0].group(2)))]
            for pm in pending:
                if pm.group(3): st.append((os.path.basename(pm.group(3)), int(pm.group(4))))
            stack = st; pending = []
        if any(f[0] == "rt_core.cuh" and lo <= f[1] <= hi for f in stack):
            n += 1
            print(f"{n:4d} {' <- '.join(f'{f[0][:7]}:{f[1]}' for f in stack[:3]):48s} {l.split('*/', 1)[1].strip()[:80]}")
