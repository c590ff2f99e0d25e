"""Registers / shared memory / stack / barriers per kernel of libmtb200.so (cuobjdump -res-usage)."""
import os, re, subprocess
so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mpi_tensorflow_b200", "lib", "libmtb200.so")
out = subprocess.run(["cuobjdump", "-res-usage", so], stdout=subprocess.PIPE, text=True).stdout
rows, fn = [], None
for line in out.splitlines():
    m = re.match(r"\s*Function (\S+):", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
        fn = re.sub(r"\(anonymous namespace\)::", "", fn).split("(")[0]
        continue
    if fn and "REG:" in line:
        g = dict(re.findall(r"(\w+):(\d+)", line))
        rows.append((fn, int(g.get("REG", 0)), int(g.get("SHARED", 0)), int(g.get("STACK", 0)), int(g.get("LOCAL", 0))))
        fn = None
print(f"{'kernel':62s} {'regs':>5s} {'static smem':>11s} {'stack':>6s} {'local':>6s}   (dynamic smem is set at launch)")
for fn, reg, sh, stack, local in rows:
    print(f"{fn[:62]:62s} {reg:5d} {sh:11d} {stack:6d} {local:6d}")
