"""SASS mnemonic counts per kernel of libraygun_b200.so -> profiles/sass_counts.txt (runs without a GPU: cuobjdump -sass).
UTCHMMA = tcgen05.mma, UTMALDG / UTMASTG = TMA load / store, LDTM / STTM = tcgen05.ld / st, SYNCS = mbarrier ops,
HMMA. = legacy mma.sync, ATOM / RED / REDG = atomics (only ATOMS.* shared-memory bit ops around tcgen05.alloc are expected).

    python scripts/sass_counts.py [> profiles/sass_counts.txt]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "raygun_b200", "lib", "libraygun_b200.so")
KEYS = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMASTG", "UTMAPF", "LDTM", "STTM", "SYNCS", "HMMA.", "ATOM", "RED.", "REDG"]

sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
names = subprocess.run(["c++filt"], input="\n".join(re.findall(r"Function : (\S+)", sass)), capture_output=True, text=True).stdout.split("\n")
demangled = dict(zip(re.findall(r"Function : (\S+)", sass), names))
per = collections.OrderedDict()
cur = None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = per.setdefault(re.sub(r"\(.*", "", demangled.get(m.group(1), m.group(1))), collections.Counter())
        continue
    m = re.search(r"/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P[0-9T]+\s+)?([A-Z][A-Z0-9_.]+)", line)
    if cur is not None and m:
        op = m.group(1)
        for k in KEYS:
            if op.startswith(k) or (k == "RED." and op.startswith("RED.")):
                cur[k] += 1
print("SASS mnemonic counts per kernel of libraygun_b200.so (cuobjdump -sass; scripts/sass_counts.py).  UTCHMMA = tcgen05.mma,")
print("UTMALDG/UTMASTG = TMA load/store, LDTM/STTM = tcgen05.ld/st, SYNCS = mbarrier ops, HMMA. = legacy mma.sync (none), ATOM = ATOMS.OR /")
print("ATOMS.AND only: SHARED-memory bit ops around tcgen05.alloc / dealloc; there is no ATOMG / RED / REDG (global atomic) instruction")
print("anywhere in the library: every floating-point reduction is two-stage and ordered.  Kernels without any of these are not listed.\n")
tot = collections.Counter()
for name, c in per.items():
    tot.update(c)
    if any(c[k] for k in KEYS):
        print(f"{name:<70s} " + "  ".join(f"{k.rstrip('.')}={c[k]}" for k in KEYS if c[k]))
print(f"\nkernels in the library: {len(per)}")
print("TOTAL " + "  ".join(f"{k}={tot[k]}" for k in KEYS))
