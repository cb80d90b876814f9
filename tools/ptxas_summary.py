#!/usr/bin/env python
"""Developer tool: registers / spills / static smem per kernel of one .cu translation unit (nvcc -Xptxas -v), one line each."""
import re
import subprocess
import sys

src = sys.argv[1]
pat = sys.argv[2] if len(sys.argv) > 2 else ""
cmd = ["/usr/local/cuda/bin/nvcc", "-O3", "-std=c++20", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
       "-fmad=false", "-Xcompiler", "-fPIC", "-Xptxas", "-v", "-c", src, "-o", "/tmp/_ptxas_summary.o"] + sys.argv[3:]
out = subprocess.run(cmd, capture_output=True, text=True)
txt = out.stderr + out.stdout
if out.returncode != 0:
    print("\n".join(l for l in txt.splitlines() if "error" in l or "^" in l)[:4000])
    sys.exit(1)
name = None
for line in txt.splitlines():
    m = re.search(r"Compiling entry function '(\S+)'", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = name.replace("b200emb::(anonymous namespace)::", "").replace("(b200emb::UpdateOp)", "op").replace("void ", "")
        name = re.sub(r">\(.*", ">", name)
        spill = ""
        continue
    m = re.search(r"(\d+) bytes spill stores", line)
    if m:
        spill = m.group(1)
    m = re.search(r"Used (\d+) registers", line)
    if m and name and pat in name:
        print("%-70s regs %3s spill %s" % (name[:70], m.group(1), spill))
