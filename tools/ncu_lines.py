"""Warp-stall samples of one captured launch aggregated per SOURCE LINE: ncu's source page gives samples per SASS
address, `nvdisasm --print-line-info` of the shipped cubin gives the source line of every SASS offset (the library is
built with -lineinfo). The .so must be the build that was profiled.
  python tools/ncu_lines.py file.ncu-rep <launch-id 1-based> <cubin stem, e.g. filter> <kernel name substring> [n]"""
import collections, csv, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rep, kid, stem, kname = sys.argv[1:5]
n = int(sys.argv[5]) if len(sys.argv) > 5 else 25
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "pyrapose_b200", "_lib", "libpyrapose_b200.so")], cwd=tmp,
               capture_output=True)
cub = [f for f in os.listdir(tmp) if f.startswith(stem + ".")][0]
dis = subprocess.run(["nvdisasm", "--print-line-info", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout
line_of, cur, active = {}, None, False
for l in dis.splitlines():
    if l.startswith("\t.section\t.text."):
        active = kname in l
    if not active:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*);", l)
    if m:
        line_of[int(m.group(1), 16)] = cur
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", ":::" + kid], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
hdr = rows[heads[0]]
i_s, i_ex = hdr.index("# Samples"), hdr.index("Instructions Executed")
stalls = [(i, h[6:]) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = [r for r in rows[heads[0] + 1:] if len(r) == len(hdr) and r[i_s].isdigit()]
base = int(data[0][0], 16)
per = collections.defaultdict(lambda: [0, 0, collections.Counter()])
for r in data:
    key = line_of.get(int(r[0], 16) - base, ("?", 0))
    per[key][0] += int(r[i_s])
    per[key][1] += int(r[i_ex])
    for i, h in stalls:
        per[key][2][h] += int(r[i] or 0)
tot = sum(v[0] for v in per.values())
print(rows[0][1][:100], "| samples", tot, "| warp instructions", sum(v[1] for v in per.values()))
src_cache = {}
for key, (s, ex, st) in sorted(per.items(), key=lambda kv: -kv[1][0])[:n]:
    f, ln = key
    path = os.path.join(ROOT, "pyrapose_b200", "csrc", f)
    if f not in src_cache and os.path.exists(path):
        src_cache[f] = open(path).read().splitlines()
    text = src_cache.get(f, [""] * (ln + 1))[ln - 1].strip() if ln else ""
    print("%5.1f%% %9d instr  %s:%d  %-70s %s" % (100.0 * s / tot, ex, f, ln, text[:70], st.most_common(2)))
