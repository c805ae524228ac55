#!/usr/bin/env python3
"""SASS census of whisper_burn_b200/libwq4.so: per kernel, how many tcgen05 / TMEM / TMA / bulk-copy / mixed-precision FMA
instructions the sm_100a binary holds (cuobjdump -sass; mnemonics from the B200 profiling guide).
Usage: python scripts/sass_census.py > profiles/r02_sass_census.txt"""
import collections, re, subprocess, sys
from pathlib import Path
so = Path(__file__).resolve().parent.parent / "whisper_burn_b200" / "libwq4.so"
out = subprocess.run(["cuobjdump", "-sass", str(so)], capture_output=True, text=True, check=True).stdout
pats = ["UTCHMMA", "UTCQMMA", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTCBAR", "SYNCS", "FHFMA", "HMMA", "LDGSTS", "MUFU.EX2", "ACQBULK", "UCGABAR"]
per = collections.OrderedDict()
cur = None
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = m.group(1); per[cur] = collections.Counter(); continue
    if cur is None: continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P[T\d]+\s+)?([A-Z0-9_.]+)", line)
    if m:
        per[cur]["_all"] += 1
        op = m.group(1)
        for p in pats:
            if op.startswith(p): per[cur][p] += 1
def demangle(n):
    try: return subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
    except Exception: return n
tot = collections.Counter()
print(f"# cuobjdump -sass {so.name} (sm_100a): instruction census per kernel; columns = SASS mnemonic prefixes")
print("kernel".ljust(86), "instr".rjust(7), " ".join(p.rjust(8) for p in pats))
for k, c in per.items():
    name = demangle(k).replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
    name = re.sub(r"^void ", "", name)
    name = name[:name.rfind("(")] if "(" in name else name
    name = name[:84]
    print(name.ljust(86), str(c["_all"]).rjust(7), " ".join(str(c[p] or ".").rjust(8) for p in pats))
    tot.update(c)
print("TOTAL".ljust(86), str(tot["_all"]).rjust(7), " ".join(str(tot[p]).rjust(8) for p in pats))
