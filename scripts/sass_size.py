"""Instruction count per kernel of csrc/libb200rt.so (cuobjdump -sass): code size matters here, the
render kernels stall on instruction fetch when the inlined traversal copies outgrow the I-cache."""
import re, subprocess, sys, os
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(__file__), "..", "python-raytracer_b200", "csrc", "libb200rt.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
name, cnt = None, {}
for l in txt.splitlines():
    m = re.search(r'Function : (\S+)', l)
    if m:
        name = m.group(1); cnt[name] = 0; continue
    if name and re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', l):
        cnt[name] += 1
for k, v in sorted(cnt.items(), key=lambda x: x[1]):
    print(f"{v:8d} instr {v*16/1024:8.1f} KiB  {k[:100]}")
