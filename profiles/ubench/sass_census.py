"""Per-kernel SASS opcode census of the built library (runs on the CPU box: cuobjdump only).
  python profiles/ubench/sass_census.py > profiles/r03_sass_census.txt
UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA tensor loads / stores, UBLKCP = bulk copies,
UTCBAR = tcgen05.commit, HMMA = legacy mma.sync (must be 0)."""
import collections
import datetime
import re
import subprocess
import sys

LIB = sys.argv[1] if len(sys.argv) > 1 else "flow-q-learning_b200/lib/libfqlpop.so"
OPS = ("UTCHMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "HMMA")
sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
demangle = lambda names: subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.split("\n")
counts, order, cur = {}, [], None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1)
        counts[cur] = collections.Counter()
        order.append(cur)
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and cur:
        op = m.group(1).split(".")[0]
        counts[cur]["instr"] += 1
        if op in OPS:
            counts[cur][op] += 1
names = dict(zip(order, demangle(order)))
print(f"cuobjdump -sass {LIB}  ({datetime.date.today()})  -- python profiles/ubench/sass_census.py")
print(" ".join(f"{o:>8}" for o in OPS) + "    instr  kernel")
key = lambda f: (-counts[f]["UTCHMMA"], -counts[f]["instr"])
for f in sorted(order, key=key):
    n = re.sub(r"fqlpop::\(anonymous namespace\)::", "", names[f])
    n = re.sub(r"\(.*", "", n) if counts[f]["UTCHMMA"] else n
    print(" ".join(f"{counts[f][o]:>8}" for o in OPS) + f"  {counts[f]['instr']:>7}  {n[:110]}")
