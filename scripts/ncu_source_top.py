"""Rank source lines of an `ncu --page source --csv --print-source cuda,sass` export by warp-stall samples."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr, cur, out = None, "", []
for r in rows:
    if r and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0] not in ("", "-"):
        out.append((cur, r))
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[6]) for _, r in out)
print("total samples", tot)
out.sort(key=lambda fr: -int(fr[1][6]))
for f, r in out[:top]:
    st = sorted(((int(r[i]), hdr[i]) for i in stall), reverse=True)[:3]
    print(f"{f[:14]:14s} {r[0]:>4s} {100 * int(r[6]) / tot:5.1f}% inst={r[7]:>8s}",
          " ".join(f"{h[6:]}={v}" for v, h in st), "|", r[1].strip()[:80])
