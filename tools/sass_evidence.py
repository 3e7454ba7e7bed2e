"""SASS evidence for the Blackwell-native kernels: counts of the tcgen05 / TMEM / TMA mnemonics per kernel of the shipped
library (cuobjdump -sass), written to profiles/.   python tools/sass_evidence.py [out.txt]
UTC*MMA = tcgen05.mma, UTCBAR = tcgen05.commit, LDTM/STTM = tcgen05.ld/st, UBLKCP / UTMALDG = cp.async.bulk (tensor),
HMMA = legacy mma.sync (must be 0 everywhere), SYNCS = mbarrier operations."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "deep-learning-at-scale_b200", "lib", "libdlas_b200.so")
PATS = ["UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTCATOMSWS", "UBLKCP", "UTMALDG", "UTMASTG", "HMMA", "SYNCS", "STG.E.64.STRONG.SYS",
        "LDG.E.64.STRONG.SYS", "STAS", "UCGABAR"]


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_sass_evidence.txt")
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels, cur = {}, None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = {p: 0 for p in PATS}
            kernels[cur]["_instr"] = 0
            continue
        if cur and re.search(r"/\*[0-9a-f]{4,}\*/", line):
            kernels[cur]["_instr"] += 1
            for p in PATS:
                if re.search(r"\b" + re.escape(p), line):
                    kernels[cur][p] += 1
    demangle = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    lines = ["# cuobjdump -sass deep-learning-at-scale_b200/lib/libdlas_b200.so | per-kernel mnemonic counts (tools/sass_evidence.py)",
             "# arch: " + ", ".join(sorted(set(re.findall(r"arch = (sm_\w+)", sass)))), ""]
    for (k, c), name in zip(kernels.items(), demangle):
        hits = " ".join("%s=%d" % (p, c[p]) for p in PATS if c[p])
        lines.append("%-6d instr  %-70s %s" % (c["_instr"], name[:70], hits or "(CUDA cores only)"))
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
