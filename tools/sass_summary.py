"""Per-kernel counts of the SASS mnemonics that show a Blackwell-native kernel (tcgen05.mma -> UTC*MMA, tcgen05.ld /
st -> LDTM / STTM, TMA -> UTMALDG, mbarrier -> SYNCS, peer / system-scope memory ops), from cuobjdump.

    python tools/sass_summary.py [lib.so] > profiles/sass/kernels_sass_summary.txt
"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "smdistributed_modelparallel_b200/lib/libsmpb200_kernels.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
PAT = re.compile(r"\b(UTC[A-Z]*MMA[.\w]*|UTCBAR[.\w]*|LDTM[.\w]*|STTM[.\w]*|UTMALDG[.\w]*|UTMASTG[.\w]*|UBLKCP[.\w]*|"
                 r"SYNCS\.[.\w]*|HMMA[.\w]*|MUFU\.EX2|MEMBAR\.[.\w]*|LDG\.E[.\w]*\.SYS|STG\.E[.\w]*\.SYS|ATOMG[.\w]*|"
                 r"UCGABAR[.\w]*|FENCE[.\w]*)")
print("SASS evidence: cuobjdump -sass", lib, "(sm_100a).")
print("Per kernel: counts of tensor-core (UTC*MMA = tcgen05.mma; .2CTA forms are cta_group::2), tensor-memory (LDTM / STTM =")
print("tcgen05.ld / st), TMA (UTMALDG), mbarrier (SYNCS), cluster barrier and system-scope / peer-memory instructions.")
print("HMMA (legacy mma.sync) must not appear.\n")
cur, counts, n = None, None, 0
def flush():
    if cur is not None and counts:
        print(f"{cur}  [{n} instructions]")
        for k, v in sorted(counts.items()):
            print(f"{v:9d}  {k}")
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        flush()
        cur, counts, n = m.group(1), collections.Counter(), 0
        continue
    if cur is None or "/*" not in line:
        continue
    body = line.split("*/", 1)[-1]
    if re.search(r"^\s+[@!A-Z]", body):
        n += 1
    for t in PAT.findall(body):
        counts[t] += 1
flush()
