"""Source-line profile of like_batch_kernel<FastImpl>: merges an ncu --set full capture (per-SASS-instruction executed
counts, `--import-source on`) with nvdisasm line info of the object file the profiled library was linked from.
Usage: python scripts/line_profile.py <prof.ncu-rep> [units_per_launch] [object-stem] [kernel-substring ...]
(run before rebuilding the library; defaults: like_fast, "like_batch_kernel" "FastImpl")"""
import collections, csv, glob, os, re, subprocess, sys, tempfile

rep = sys.argv[1]
evals = float(sys.argv[2]) if len(sys.argv) > 2 else 4096 * 10000
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
obj = os.path.join(root, "hddm_b200", "csrc", "_obj", (sys.argv[3] if len(sys.argv) > 3 else "like_fast") + ".o")
pats = sys.argv[4:] if len(sys.argv) > 4 else ["like_batch_kernel", "FastImpl"]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", obj], cwd=tmp, check=True, capture_output=True)
cubin = glob.glob(os.path.join(tmp, "*.cubin"))[0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
# locate the batch kernel section
start = next(i for i, l in enumerate(dis) if l.startswith("//---") and all(p in l for p in pats))
end = next((i for i in range(start + 1, len(dis)) if dis[i].startswith("//---")), len(dis))
lines = []  # per instruction: (file, line, outermost inlined-at line)
cur = ("?", 0)
for l in dis[start:end]:
    m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/", l):
        op = re.sub(r"^\s*/\*[0-9a-f]+\*/\s*", "", l).split(";")[0].strip()
        lines.append((cur, op))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[1]
iS, iE = hdr.index("Source"), hdr.index("Instructions Executed")
counts = []
for r in rows[2:]:
    try:
        counts.append((int(r[iE]), r[iS].strip()))
    except (ValueError, IndexError):
        pass
print(f"# disasm instructions {len(lines)}, profiled instructions {len(counts)}")
n = min(len(lines), len(counts))
evalw = evals / 32
by_line = collections.Counter()
by_line_fp64 = collections.Counter()
for (loc, op), (c, s) in zip(lines[:n], counts[:n]):
    by_line[loc] += c
    if re.match(r"(@!?U?P\d+\s+)?(DFMA|DMUL|DADD|DSETP)", s):
        by_line_fp64[loc] += c
tot = sum(by_line.values())
print(f"# total warp instructions per eval: {tot / evalw:.0f}")
for loc, c in by_line.most_common(60):
    print(f"{c / evalw:7.1f} ({by_line_fp64[loc] / evalw:6.1f} fp64)  {loc[0]}:{loc[1]}")
