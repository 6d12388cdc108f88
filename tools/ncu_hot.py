"""Top stall locations of one captured launch: python tools/ncu_hot.py file.ncu-rep <launch-id 1-based> [n]"""
import csv, subprocess, sys
rep, kid = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", ":::" + kid], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
i_src, i_s, i_ex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stalls = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = rows[2:]
tot = sum(int(r[i_s]) for r in data)
print("total samples", tot, " instructions", sum(int(r[i_ex]) for r in data))
agg = sorted(((sum(int(r[i] or 0) for r in data), h) for i, h in stalls), reverse=True)[:8]
print(" ".join("%s=%d" % (h[6:], v) for v, h in agg))
for k, r in enumerate(data):
    r.append(k)
for r in sorted(data, key=lambda r: -int(r[i_s]))[:n]:
    st = sorted([(int(r[i] or 0), h[6:]) for i, h in stalls], reverse=True)[:2]
    print(str(r[-1]).rjust(5), r[i_s].rjust(6), r[i_ex].rjust(9), r[i_src].strip()[:64].ljust(64), st)
