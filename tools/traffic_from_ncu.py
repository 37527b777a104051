"""profiles/traffic.json from an ncu --set full capture of ONE mapping view (all sb_* kernels of the view, in order):
per stage, the measured DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) and executed warp instructions
(smsp__inst_executed.sum) summed over the stage's launches.  bench.py reads the file for roofline.traffic / .issue.

Usage: python tools/traffic_from_ncu.py REPORT.ncu-rep C2 [CSRC_HASH] > profiles/traffic.json"""
import csv
import importlib.util
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _stamp():
    """git SHA of HEAD and the content hash of csrc/ at capture-summary time (the capture must be of this build)."""
    spec = importlib.util.spec_from_file_location("sb200_build", os.path.join(ROOT, "vings-mono_b200", "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    sha = subprocess.run(["git", "-C", ROOT, "rev-parse", "HEAD"], capture_output=True, text=True).stdout.strip()
    return sha, mod.csrc_hash()

# kernel name prefix -> stage (the names of include/surfel_b200.h's stage enum as bench.py spells them)
STAGE_OF = [
    ("sb_activate", "activate"), ("sb_preprocess_backward", "preprocess_bwd"), ("sb_preprocess", "preprocess"),
    ("sb_sort_hist", "surfel_sort"), ("sb_sort_scan", "surfel_sort"), ("sb_onesweep_kernel<8>", "surfel_sort"),
    ("sb_onesweep_kernel<16>", "sort"), ("sb_pair_offsets", "duplicate"), ("sb_duplicate", "duplicate"),
    ("sb_tile_scan", "ranges"), ("sb_blend_forward", "blend_fwd"), ("sb_blend_backward", "blend_bwd"),
    ("sb_loss", "map_loss"), ("sb_ssim", "map_loss"), ("sb_geom_loss", "map_loss"), ("sb_surfel_update", "update"),
    ("sb_view_update", "update"), ("sb_dp_", "update"),
    ("sb_adam", "adam"),
]
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "inst": 1.0}


def main():
    rep, cfg = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    ki = hdr.index("Kernel Name")
    cols = {n: hdr.index(n) for n in ("dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum")}
    traffic, inst = {}, {}
    for r in body:
        name = r[ki].replace("void ", "")
        stage = next((s for p, s in STAGE_OF if name.startswith(p)), None)
        if stage is None:
            continue
        val = {n: float(r[i]) * UNIT.get(units[i], 1.0) for n, i in cols.items()}
        traffic[stage] = traffic.get(stage, 0.0) + val["dram__bytes_read.sum"] + val["dram__bytes_write.sum"]
        inst[stage] = inst.get(stage, 0.0) + val["smsp__inst_executed.sum"]
    sha, chash = _stamp()
    if len(sys.argv) > 3:   # the hash the capture was made with (written next to the report on the GPU box)
        chash = sys.argv[3]
    out = {"_git_sha": sha, "_csrc_hash": chash,
           "_source": "%s: one ncu --set full capture of every kernel of one %s mapping view (dram__bytes_read.sum + "
                      "dram__bytes_write.sum and smsp__inst_executed.sum, summed over the launches of a stage); "
                      "made by tools/traffic_from_ncu.py" % (rep, cfg),
           cfg: traffic, cfg + "_warp_inst": inst}
    json.dump(out, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
