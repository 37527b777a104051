"""Per-source-line profile of one kernel from an ncu report (built with -lineinfo, captured with --import-source on).

ncu's CSV source page is per SASS instruction; nvdisasm -g gives the line of every SASS offset of the same
cubin.  This joins the two (by instruction offset within the kernel) and prints, per source line, the share
of executed warp instructions and of stall samples, plus the top stall reasons.

Usage: python tools/ncu_lines.py REPORT.ncu-rep KERNEL_REGEX [path/to/lib.so] [--min 0.5] [--sass]
"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT_SO = os.path.join(ROOT, "vings-mono_b200", "diff_surfel_rasterization", "libsurfel_b200.so")


def sass_lines(so, kernel_re):
    """offset -> (file, line) for the first kernel whose mangled name matches kernel_re."""
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, check=True, capture_output=True)
    for f in sorted(os.listdir(tmp)):
        if "sm_100" not in f:
            continue
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        cur, line, out = None, None, {}
        for ln in txt.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", ln)
            if m:
                if out:
                    return out
                cur = m.group(1) if re.search(kernel_re, m.group(1)) else None
                continue
            if cur is None:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                line = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]+)\*/\s+(.*?);", ln)
            if m:
                out[int(m.group(1), 16)] = (line, m.group(2).strip())
        if out:
            return out
    raise SystemExit("kernel %s not found in %s" % (kernel_re, so))


def main():
    argv = sys.argv[1:]
    if "--min" in argv:
        del argv[argv.index("--min") + 1]
    args = [a for a in argv if not a.startswith("--")]
    rep, kre = args[0], args[1]
    so = args[2] if len(args) > 2 else DEFAULT_SO
    thr = float(sys.argv[sys.argv.index("--min") + 1]) if "--min" in sys.argv else 0.5
    r = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre],
                       capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(r)))
    h = next(i for i, x in enumerate(rows) if x and x[0] == "Address")
    hdr = rows[h]
    col = {n: i for i, n in enumerate(hdr)}
    body = []
    for x in rows[h + 1:]:
        if len(x) < len(hdr) or x[0] == "Address":
            break  # only the first matching launch
        body.append(x)
    base = int(body[0][0], 16)
    lines = sass_lines(so, kre)
    stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
    per = defaultdict(lambda: [0.0, 0.0, defaultdict(float), 0.0])
    tot_i = tot_s = 0.0
    sass_rows = []
    for x in body:
        off = int(x[0], 16) - base
        loc, _ = lines.get(off, ((("?", 0)), ""))
        inst = float(x[col["Instructions Executed"]] or 0)
        samp = float(x[col["# Samples"]] or 0)
        thr_i = float(x[col["Thread Instructions Executed"]] or 0)
        e = per[loc]
        e[0] += inst
        e[1] += samp
        e[3] += thr_i
        for s in stall_cols:
            v = float(x[col[s]] or 0)
            if v:
                e[2][s] += v
        tot_i += inst
        tot_s += samp
        sass_rows.append((off, loc, x[col["Source"]], inst, samp))
    print("kernel %s: %.4g warp instructions, %d samples" % (kre, tot_i, tot_s))
    srcs = {}
    for loc, e in sorted(per.items(), key=lambda kv: (kv[0] is None, kv[0])):
        if loc is None:
            continue
        pi, ps = 100 * e[0] / tot_i, 100 * e[1] / max(tot_s, 1)
        if pi < thr and ps < thr:
            continue
        f, l = loc
        if f not in srcs:
            p = [os.path.join(dp, f) for dp, _, fs in os.walk(ROOT) if f in fs]
            srcs[f] = open(p[0]).read().splitlines() if p else []
        text = srcs[f][l - 1].strip() if 0 < l <= len(srcs[f]) else ""
        top = sorted(e[2].items(), key=lambda kv: -kv[1])[:3]
        stalls = " ".join("%s:%.0f%%" % (k[6:], 100 * v / max(e[1], 1)) for k, v in top)
        act = e[3] / max(e[0], 1)
        print("%-22s %5.1f%% inst %5.1f%% samp  act %4.1f | %-70s | %s" % ("%s:%d" % (f, l), pi, ps, act, text[:70], stalls))
    if "--sass" in sys.argv:
        for off, loc, s, inst, samp in sass_rows:
            if 100 * inst / tot_i >= thr / 4 or 100 * samp / max(tot_s, 1) >= thr / 4:
                print("%05x %-22s %5.2f%% %5.2f%% %s" % (off, "%s:%d" % loc if loc else "?", 100 * inst / tot_i,
                                                        100 * samp / max(tot_s, 1), s))


if __name__ == "__main__":
    main()
