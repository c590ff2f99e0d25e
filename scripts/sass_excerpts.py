"""Per-kernel SASS excerpt of libmtb200.so: every tensor-core / TMA / TMEM / mbarrier / multimem / peer-memory
instruction with its address (cuobjdump -sass, encodings stripped).  Complements sass_evidence.py (counts only)."""
import os, re, subprocess
so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mpi_tensorflow_b200", "lib", "libmtb200.so")
out = subprocess.run(["cuobjdump", "-sass", so], stdout=subprocess.PIPE, text=True).stdout
keep = re.compile(r"\b(UTC\w*MMA|UTCBAR|UTCATOMSWS|LDTM|STTM|UTMALDG|UTMASTG|UTMAPF|UTMACCTL|SYNCS|LDGMC|STGMC|REDGMC|ELECT|"
                  r"ACQBULK|PREEXIT|ACQSHMINIT|FENCE\.VIEW\.ASYNC\S*|LD\.E\S*\.SYS\S*|LDG\.E\S*\.SYS\S*|ST\.E\S*\.SYS\S*|"
                  r"STG\.E\S*\.SYS\S*|REDG\.E\S*\.SYS|UCGABAR_ARV|UCGABAR_WAIT|HMMA)\b")
fn, lines, total = None, [], 0
def flush():
    global lines
    if fn and lines:
        print(f"\n== {fn}  ({total} instructions, {len(lines)} shown)")
        print("\n".join(lines))
    lines = []
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        flush()
        fn = subprocess.run(["c++filt", m.group(1)], stdout=subprocess.PIPE, text=True).stdout.strip()
        fn = re.sub(r"\(anonymous namespace\)::", "", fn).split("(")[0]
        total = 0
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m:
        total += 1
        if keep.search(m.group(2)):
            lines.append(f"  /*{m.group(1)}*/  {m.group(2).strip()}")
flush()
