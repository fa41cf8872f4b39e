"""Per-kernel SASS evidence: counts of the mnemonics that prove (or disprove) a Blackwell-native kernel.

    python scripts/sass_summary.py [ddim_cold_b200/libddimcold.so] > profiles/sass_summary.txt

UTC*MMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / st, UTMALDG / UTMASTG = TMA tensor load / store, UTCBAR =
tcgen05.commit, HMMA = legacy mma.sync, LDGSTS = cp.async, MUFU = special-function unit.
"""
import collections
import re
import subprocess
import sys

so = sys.argv[1] if len(sys.argv) > 1 else "ddim_cold_b200/libddimcold.so"
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
pats = collections.OrderedDict([("UTC*MMA", r"\bUTC[A-Z]*MMA"), ("LDTM", r"\bLDTM"), ("STTM", r"\bSTTM"), ("UTMALDG", r"\bUTMALDG"),
                                ("UTMASTG", r"\bUTMASTG"), ("UTCBAR", r"\bUTCBAR"), ("HMMA", r"\bHMMA"), ("LDGSTS", r"\bLDGSTS"),
                                ("MUFU", r"\bMUFU"), ("FFMA", r"\bFFMA")])
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
rows, cur, counts = [], None, None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        if cur:
            rows.append((cur, counts))
        cur, counts = m.group(1), collections.Counter()
        continue
    if cur:
        for k, p in pats.items():
            if re.search(p, line):
                counts[k] += 1
if cur:
    rows.append((cur, counts))
print(f"# cuobjdump -sass {so}  ({len(rows)} functions, sm_100a)")
print("# " + "  ".join(f"{k:>8s}" for k in pats) + "  kernel")
tot = collections.Counter()
for name, c in sorted(rows, key=lambda r: demangle(r[0])):
    tot.update(c)
    short = re.sub(r"\(.*", "", demangle(name)).replace("void ", "").replace("ddc::", "")
    print("  " + "  ".join(f"{c[k]:8d}" for k in pats) + "  " + short)
print("# " + "  ".join(f"{tot[k]:8d}" for k in pats) + "  TOTAL")
