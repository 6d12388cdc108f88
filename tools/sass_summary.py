"""Opcode census of the shipped library: `cuobjdump -sass` of pyrapose_b200/_lib/libpyrapose_b200.so, per kernel the
count of the SASS mnemonics that prove what the kernel runs on (tcgen05 = UTCHMMA / UTCBAR / LDTM, TMA = UTMALDG /
UTMASTG / UBLKCP, legacy tensor path = HMMA, which must stay at 0). Usage: python tools/sass_summary.py [out.txt]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "pyrapose_b200", "_lib", "libpyrapose_b200.so")
WATCH = ["UTCHMMA", "UTCHMMA.2CTA", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "HMMA", "FADD2", "HMNMX2",
         "DFMA", "DMUL", "DADD", "MUFU.RCP64H", "REDG", "ATOMG", "RED", "LDG.E.128", "STG.E.128", "SHFL", "VOTE", "BAR"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    out = [o.replace("(anonymous namespace)::", "").replace("ppb::", "") for o in out]
    return [re.sub(r"\(.*$", "", re.sub(r"^void ", "", o)) for o in out]


def main(dst):
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    per, cur, arch = collections.OrderedDict(), None, set()
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = per.setdefault(m.group(1), collections.Counter())
            continue
        m = re.match(r"\s*arch = (\S+)", line)
        if m:
            arch.add(m.group(1))
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and cur is not None:
            op = m.group(1)
            cur["_total"] += 1
            for w in WATCH:
                if op == w or op.startswith(w + "."):
                    cur[w] += 1
            if op.startswith("UTCHMMA") and ".2CTA" in op:
                cur["UTCHMMA.2CTA"] += 1
    names = demangle(list(per))
    cols = [w for w in WATCH if any(c[w] for c in per.values()) or w == "HMMA"]
    with open(dst, "w") as f:
        f.write("# cuobjdump -sass pyrapose_b200/_lib/libpyrapose_b200.so ; arch %s ; %d kernels\n" % (sorted(arch), len(per)))
        f.write("# tcgen05: UTCHMMA (MMA issue; .2CTA = cta_group::2), UTCBAR (tcgen05.commit), LDTM (tcgen05.ld); TMA: UTMALDG /\n"
                "# UTMASTG (tensor load / store), UBLKCP (cp.async.bulk); HMMA (mma.sync) must be 0 everywhere\n")
        f.write("%-46s %7s " % ("kernel", "instrs") + " ".join("%*s" % (max(6, len(c)), c) for c in cols) + "\n")
        tot = collections.Counter()
        for n, c in sorted(zip(names, per.values()), key=lambda t: t[0]):
            f.write("%-46s %7d " % (n[:46], c["_total"]) + " ".join("%*d" % (max(6, len(k)), c[k]) for k in cols) + "\n")
            tot.update(c)
        f.write("%-46s %7d " % ("TOTAL", tot["_total"]) + " ".join("%*d" % (max(6, len(k)), tot[k]) for k in cols) + "\n")
    print(open(dst).read())


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_summary.txt"))
