"""Top stall locations of one captured launch: python tools/ncu_hot.py file.ncu-rep <launch-id 1-based> [n]"""
import csv, subprocess, sys
rep, kid = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", ":::" + kid], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
heads = [i for i, r in enumerate(rows) if r and r[0] in ("Address", "#")]
hdr = rows[heads[0]]
end = heads[1] - 1 if len(heads) > 1 else len(rows)
print(rows[0][:2])
i_src, i_s, i_ex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stalls = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = [r for r in rows[heads[0] + 1:end] if len(r) == len(hdr) and r[i_s].isdigit()]
tot = sum(int(r[i_s]) for r in data)
print("total samples", tot, " instructions", sum(int(r[i_ex]) for r in data))
agg = sorted(((sum(int(r[i] or 0) for r in data), h) for i, h in stalls), reverse=True)[:8]
print(" ".join("%s=%d" % (h[6:], v) for v, h in agg))
for k, r in enumerate(data):
    r.append(k)
for r in sorted(data, key=lambda r: -int(r[i_s]))[:n]:
    st = sorted([(int(r[i] or 0), h[6:]) for i, h in stalls], reverse=True)[:2]
    print(str(r[-1]).rjust(5), r[i_s].rjust(6), r[i_ex].rjust(9), r[i_src].strip()[:64].ljust(64), st)
