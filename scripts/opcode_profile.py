"""Executed warp instructions per unit of work by opcode, from an ncu --set full capture (--import-source on).
    python scripts/opcode_profile.py <prof.ncu-rep> <units_per_launch>"""
import collections, csv, re, subprocess, sys

rep, units = sys.argv[1], float(sys.argv[2])
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[1]
iS, iE = hdr.index("Source"), hdr.index("Instructions Executed")
by = collections.Counter()
tot = 0
for r in rows[2:]:
    try:
        c = int(r[iE])
    except (ValueError, IndexError):
        continue
    s = re.sub(r"^(@!?U?P\d+\s+)", "", r[iS].strip())
    op = s.split()[0].split(".")[0] if s else "?"
    full = s.split()[0] if s else "?"
    key = op if op not in ("LDS", "LDL", "STL", "LDG", "LD", "ST", "STS", "LDC", "ISETP", "IMAD", "MUFU") else ".".join(full.split(".")[:2])
    by[key] += c
    tot += c
w = units / 32
print(f"# total warp instructions per unit: {tot / w:.1f}")
for k, c in by.most_common(45):
    print(f"{c / w:8.2f}  {k}")
