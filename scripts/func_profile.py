"""Per-function split of an ncu capture of a likelihood kernel: sums the source page's per-instruction counters over the
address ranges of the kernel body and of each out-of-line device function in the same section.
Usage: python scripts/func_profile.py <prof.ncu-rep> <object.o> <kernel-substring>   (object = the build the capture ran)"""
import csv, glob, os, re, subprocess, sys, tempfile, collections

rep, obj, pat = os.path.abspath(sys.argv[1]), os.path.abspath(sys.argv[2]), sys.argv[3]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd=tmp, check=True, capture_output=True)
cubin = glob.glob(os.path.join(tmp, "*.cubin"))[0]
sym = subprocess.run(["readelf", "-s", "-W", cubin], capture_output=True, text=True).stdout.splitlines()
sec = subprocess.run(["readelf", "-S", "-W", cubin], capture_output=True, text=True).stdout.splitlines()
secidx = None
for l in sec:
    m = re.match(r"\s*\[\s*(\d+)\]\s+(\.text\.\S+)", l)
    if m and pat in m.group(2):
        secidx = int(m.group(1))
funcs = []
for l in sym:
    f = l.split()
    if len(f) >= 8 and f[3] == "FUNC" and f[6] == str(secidx):
        funcs.append((int(f[1], 16), int(f[2]), f[7]))
funcs.sort()
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[1]
ia, ie, isamp, ino, ibar, ith = (hdr.index(k) for k in ("Address", "Instructions Executed", "# Samples", "stall_no_inst", "stall_barrier", "Thread Instructions Executed"))
base = int(rows[2][ia], 16)
agg = collections.defaultdict(lambda: [0, 0, 0, 0, 0])
for r in rows[2:]:
    try:
        off = int(r[ia], 16) - base
    except (ValueError, IndexError):
        continue
    name = "?"
    for o, sz, n in funcs:
        if o <= off < o + sz:
            name = n
    a = agg[name]
    a[0] += int(r[ie]); a[1] += int(r[isamp]); a[2] += int(r[ino]); a[3] += int(r[ibar]); a[4] += int(r[ith])
tot = sum(a[1] for a in agg.values()) or 1
print(f"{'function':70s} {'warp-instr':>12s} {'lanes':>6s} {'samples%':>8s} {'no_inst%':>8s} {'barrier%':>8s}")
for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    short = n.split("$")[-1][-70:]
    print(f"{short:70s} {a[0]:12d} {a[4] / max(a[0], 1):6.1f} {100 * a[1] / tot:8.1f} {100 * a[2] / tot:8.1f} {100 * a[3] / tot:8.1f}")
