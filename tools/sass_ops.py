"""Counts of the Blackwell-specific SASS mnemonics per kernel of libsurfel_b200.so (cuobjdump -sass), the evidence
behind DESIGN.md's "cp.async / TMA bulk copies / mbarrier / redux / match" statements.
Usage: python tools/sass_ops.py > profiles/r02_sass_blackwell_ops.txt"""
import os
import re
import subprocess
import sys
from collections import Counter, OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "vings-mono_b200", "diff_surfel_rasterization", "libsurfel_b200.so")
OPS = ["LDGSTS", "ARRIVES", "UBLKCP", "UTMALDG", "SYNCS", "REDUX", "MATCH", "RED", "REDG", "ATOMG", "ATOMS", "SHFL", "VOTE", "MUFU", "LDS", "LDG",
       "STG", "LDL", "STL", "UTCHMMA", "UTCQMMA", "HMMA"]


def main():
    so = sys.argv[1] if len(sys.argv) > 1 else SO
    txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
    archs = sorted(set(re.findall(r"arch = (sm_\w+)", txt)))
    kernels = OrderedDict()
    cur = None
    for line in txt.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
            kernels.setdefault(cur, Counter())
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur:
            kernels[cur][m.group(1)] += 1
            kernels[cur]["_total"] += 1
    print("# %s" % os.path.relpath(so, ROOT))
    print("# cubin architectures: %s   (cuobjdump -sass | per-kernel mnemonic counts; LDGSTS = cp.async 16 B global -> shared,"
          % ", ".join(archs))
    print("#  ARRIVES = cp.async.mbarrier.arrive (LDGSTSBAR), UBLKCP = cp.async.bulk (TMA 1-D; the SB_RING_LDGSTS=0 build of the ring),")
    print("#  SYNCS = mbarrier ops, REDUX = warp redux.sync, MATCH = match.any; no UTMALDG / UTC*MMA / HMMA: no tensor tiles, no MMA)")
    used = [o for o in OPS if any(k[o] for k in kernels.values())]
    print("%-44s %7s " % ("kernel", "instr") + " ".join("%7s" % o for o in used))
    for name, c in kernels.items():
        print("%-44s %7d " % (name[:44], c["_total"]) + " ".join("%7d" % c[o] for o in used))
    absent = [o for o in OPS if o not in used]
    print("# absent everywhere: %s" % ", ".join(absent))


if __name__ == "__main__":
    main()
