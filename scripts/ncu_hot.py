"""Attribute ncu warp-stall samples (source page of a .ncu-rep) to source lines WITH their inline stacks.
usage: ncu_hot.py report.ncu-rep kernel_pattern [libb200rt.so]   -- joins ncu's per-SASS rows with `nvdisasm -gi`."""
import collections, csv, io, os, re, subprocess, sys, tempfile
rep, pat = sys.argv[1], sys.argv[2]
so = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "python-raytracer_b200", "csrc", "libb200rt.so")
d = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=d, capture_output=True)
cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
txt = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
on, stack, instrs = False, [], []
pending = []
for l in txt.splitlines():
    if l.startswith("//-----") and ".text." in l:
        on = pat in l; continue
    if not on: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)(?: inlined at "([^"]+)", line (\d+))?', l)
    if m:
        pending.append(m); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', l):
        if pending:
            st = [(os.path.basename(pending[0].group(1)), int(pending[0].group(2)))]
            for pm in pending:
                if pm.group(3): st.append((os.path.basename(pm.group(3)), int(pm.group(4))))
            stack = st; pending = []
        instrs.append((l.split("*/", 1)[1].strip(), stack))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hdr_i]; body = rows[hdr_i + 1:]
si, ei = h.index("# Samples"), h.index("Instructions Executed")
stall_cols = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
assert len(body) == len(instrs), (len(body), len(instrs))
tot = sum(int(r[si]) for r in body); tot_e = sum(int(r[ei]) for r in body)
print(f"samples {tot}, warp instructions executed {tot_e}")
def agg(keyf, title, top=25):
    a, e = collections.Counter(), collections.Counter()
    for r, (ins, st) in zip(body, instrs):
        k = keyf(ins, st); a[k] += int(r[si]); e[k] += int(r[ei])
    print(f"\n== {title}")
    for k, v in a.most_common(top):
        print(f"{100*v/tot:6.2f}% samples {100*e[k]/tot_e:6.2f}% instr  {k}")
agg(lambda ins, st: st[0] if st else None, "innermost line")
agg(lambda ins, st: tuple(st[-3:-1]) if len(st) >= 3 else tuple(st), "outer frames (2 below the kernel)")
agg(lambda ins, st: next((f for f in st if f[0] == "rt_core.cuh" and f[1] > 280), st[0] if st else None), "first rt_core frame past the walker", int(os.environ.get("TOP", "25")))
stall = collections.Counter()
for r in body:
    for i in stall_cols: stall[h[i]] += int(r[i])
print("\n== stall reasons", {k: f"{100*v/tot:.1f}%" for k, v in stall.most_common(8)})
agg(lambda ins, st: ins.split()[0] if not ins.startswith("@") else ins.split()[1], "opcode", 15)
if len(sys.argv) > 4:
    col = h.index(sys.argv[4])
    a = collections.Counter()
    for r, (ins, st) in zip(body, instrs):
        a[next((f for f in st if f[0] == "rt_core.cuh" and f[1] > 280), st[0] if st else None)] += int(r[col])
    t = sum(a.values())
    print(f"\n== {sys.argv[4]} by first rt_core frame past the walker (total {t})")
    for k, v in a.most_common(12):
        print(f"{100*v/t:6.2f}%  {k}")
ti = h.index("Thread Instructions Executed")
a, e = collections.Counter(), collections.Counter()
for r, (ins, st) in zip(body, instrs):
    k = next((f for f in st if f[0] == "rt_core.cuh" and f[1] > 280), st[0] if st else None)
    a[k] += int(r[ti]); e[k] += int(r[ei])
print("\n== avg active threads per executed warp instruction, by first rt_core frame past the walker")
for k, v in e.most_common(10):
    print(f"{a[k]/max(1,v):6.2f} lanes  {100*v/tot_e:6.2f}% instr  {k}")
