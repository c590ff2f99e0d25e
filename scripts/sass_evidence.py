"""Count the Blackwell-native SASS mnemonics per kernel in libmtb200.so (cuobjdump -sass)."""
import collections, re, subprocess, os
so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mpi_tensorflow_b200", "lib", "libmtb200.so")
out = subprocess.run(["cuobjdump", "-sass", so], stdout=subprocess.PIPE, text=True).stdout
pat = {"UTC*MMA (tcgen05.mma)": r"\bUTC\w*MMA\b", "LDTM (tcgen05.ld)": r"\bLDTM\b", "UTMALDG (TMA load)": r"\bUTMALDG\b",
       "UTCBAR (tcgen05.commit)": r"\bUTCBAR\b", "SYNCS (mbarrier)": r"\bSYNCS\b", "multimem": r"\bLDGMC\b|\bSTGMC\b|\bREDGMC\b|MULTIMEM",
       "LD.E.*.SYS (peer loads)": r"\bLD\.E\S*\.SYS\b|LDG\.E\S*\.SYS", "REDG..SYS (peer atomics)": r"\bREDG\.E\S*\.SYS\b", "HMMA (legacy mma.sync)": r"\bHMMA\b"}
fn = None
counts = collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
        fn = re.sub(r"\(anonymous namespace\)::", "", fn).split("(")[0]
        counts[fn] = collections.Counter()
        continue
    if fn:
        for k, p in pat.items():
            if re.search(p, line):
                counts[fn][k] += 1
print("arch:", re.search(r"arch = (\S+)", out).group(1))
print(f"{'kernel':58s} " + " ".join(f"{k.split()[0]:>9s}" for k in pat))
for fn, c in counts.items():
    if sum(c.values()):
        print(f"{fn[:58]:58s} " + " ".join(f"{c[k]:9d}" for k in pat))
