"""Turns `ncu --csv --metrics ...` captures of ONE launch of the solve kernel into profiles/r02_roofline_counts.json:
FP64 flops per QP (2 x DFMA + DADD + DMUL thread instructions, predicated-on) and DRAM bytes per launch.
usage: python scripts/roofline_counts.py <key>=<ncu.csv>:<qps_in_launch> ...   (key: the bench workload, "stream" / "dmpc")"""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out_path = os.path.join(ROOT, "profiles", "r02_roofline_counts.json")
out = json.load(open(out_path)) if os.path.exists(out_path) else {}
for arg in sys.argv[1:]:
    key, rest = arg.split("=")
    path, qps = rest.split(":")
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    iname, ival = hdr.index("Metric Name"), hdr.index("Metric Value")
    m = {}
    for r in rows[1:]:
        m[r[iname]] = float(r[ival].replace(",", ""))
    dfma = m["smsp__sass_thread_inst_executed_op_dfma_pred_on.sum"]
    dadd = m["smsp__sass_thread_inst_executed_op_dadd_pred_on.sum"]
    dmul = m["smsp__sass_thread_inst_executed_op_dmul_pred_on.sum"]
    flops = 2 * dfma + dadd + dmul
    out[key] = {"qps_in_launch": int(qps), "dfma": dfma, "dadd": dadd, "dmul": dmul, "fp64_flops_per_launch": flops,
                "fp64_flops_per_qp": flops / int(qps),
                "dram_bytes_per_launch": m.get("dram__bytes_read.sum", 0.0) + m.get("dram__bytes_write.sum", 0.0),
                "kernel_ns_under_ncu": m.get("gpu__time_duration.sum"), "source": "profiles/" + os.path.basename(path)}
json.dump(out, open(out_path, "w"), indent=1)
print(json.dumps(out, indent=1))
