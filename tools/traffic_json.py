"""profiles/k_match_traffic.json from an ncu capture of k_match (one launch, --set full).

usage: python tools/traffic_json.py <workload name> <capture.ncu-rep> ["when / how captured"]

bench.py copies these per-launch figures into `roofline.traffic` / `dram_fraction` /
`issue_fraction` / `lsu_fraction` -- only while dstack_b200/csrc/kernels.cuh still hashes to
`kernels_cuh_sha16`, the source the capture was taken from."""
import csv
import hashlib
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
name, rep = sys.argv[1], sys.argv[2]
note = sys.argv[3] if len(sys.argv) > 3 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(out)))
head, units, vals = rows[0], rows[1], rows[2]


def get(metric, scale_to_bytes=False):
    i = head.index(metric)
    v = float(vals[i].replace(",", ""))
    if scale_to_bytes:
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
    return v


n_sm = int(get("device__attribute_multiprocessor_count")) if "device__attribute_multiprocessor_count" in head else 148
path = os.path.join(ROOT, "profiles", "k_match_traffic.json")
doc = json.load(open(path)) if os.path.exists(path) else {}
sha = hashlib.sha256(open(os.path.join(ROOT, "dstack_b200", "csrc", "kernels.cuh"), "rb").read()).hexdigest()[:16]
if doc.get("kernels_cuh_sha16") != sha:
    doc = {k: v for k, v in doc.items() if k in ()}  # captures of another kernel source are dropped
doc["kernels_cuh_sha16"] = sha
doc["captured"] = note
doc[name] = {
    "kernel": vals[head.index("Kernel Name")],
    "dram_read_bytes": int(get("dram__bytes_read.sum", True)),
    "dram_write_bytes": int(get("dram__bytes_write.sum", True)),
    "warp_instructions": int(get("smsp__inst_executed.sum")),
    # LSU data-pipe wavefronts of all kinds (shared + global + local): ncu's triage section has the
    # per-SM average
    "lsu_wavefronts": int(get("SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts.avg") * n_sm),
    "shared_wavefronts": int(get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")),
    "duration_us_under_ncu": get("gpu__time_duration.sum") * {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}[
        units[head.index("gpu__time_duration.sum")]],
    "lsu_pipe_pct": get("l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed"),
    "issue_active_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
}
json.dump(doc, open(path, "w"), indent=1)
print(json.dumps(doc[name], indent=1))
