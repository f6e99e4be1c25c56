"""profiles/rNN_sass_summary.txt: per-kernel counts of the SASS mnemonics that show how the library moves data
(UBLKCP = 1-D bulk/TMA copies, UTMALDG = tensor-map TMA, UTCMMA/LDTM = tcgen05, SYNCS = mbarrier, LDGSTS = cp.async,
DFMA/DMUL/DADD = float64 lattice, MUFU = ex2/lg2, REDUX).

    python tools/sass_summary.py > profiles/r02_sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "espnet_b200", "libespnet_b200_ctc.so")
WANT = ["UBLKCP", "UTMALDG", "UTCMMA", "UTCHMMA", "LDTM", "HMMA", "SYNCS", "UCGABAR", "LDGSTS", "DFMA", "DMUL", "DADD", "MUFU",
        "REDUX", "SHFL", "LDG", "STG", "LDS", "STS", "ATOM", "RED"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kern, counts = None, collections.OrderedDict()
    arch = set(re.findall(r"arch = (sm_\w+)", out))
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            kern = re.sub(r"\(.*", "", kern).replace("b200::", "")
            counts[kern] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and kern:
            op = m.group(1)
            counts[kern]["total"] += 1
            for w in WANT:
                if op.startswith(w):
                    counts[kern][w] += 1
                    break
    print("library:", os.path.relpath(LIB, ROOT), "cubins:", ", ".join(sorted(arch)))
    print("%-78s %6s " % ("kernel", "insts") + " ".join("%7s" % w for w in WANT))
    for k, c in counts.items():
        print("%-78s %6d " % (k[:78], c["total"]) + " ".join("%7d" % c[w] for w in WANT))
    tot = collections.Counter()
    for c in counts.values():
        tot.update(c)
    print("%-78s %6d " % ("TOTAL", tot["total"]) + " ".join("%7d" % tot[w] for w in WANT))


if __name__ == "__main__":
    sys.exit(main())
