"""Reduce a `ncu --set full` report of one mapping iteration (tests/gpu_profile_step.py) to the per-stage figures
bench.py attaches to its roofline block: DRAM bytes moved (dram__bytes_read + dram__bytes_write) and, for the blend
kernels, the fp32 work actually EXECUTED (2 x FFMA + FMUL + FADD thread instructions, predicated on; packed FFMA2
counts as two FFMA per lane) -- beside the algorithmic figures of SURVEY 8d, which count the reference's operations.
  python tests/ncu_traffic.py gpurun_out/full.ncu-rep > profiles/r02_traffic.json      (run here, no GPU needed)"""
import csv
import io
import json
import subprocess
import sys

STAGE_OF = (("preprocess_bwd", "preprocess_bwd"), ("preprocess_kernel", "preprocess"), ("onesweep", "depth_sort_offsets"),
            ("sort_", "depth_sort_offsets"), ("coverage", "depth_sort_offsets"), ("offsets", "depth_sort_offsets"),
            ("l1_", "binning"), ("render_fwd", "render_fwd"), ("render_bwd", "render_bwd"))


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        try:
            return float(r[col[name]].replace(",", ""))
        except (KeyError, ValueError):
            return 0.0

    units = rows[1]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    agg = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        stage = next((s for k, s in STAGE_OF if k in name), None)
        if stage is None:
            continue
        a = agg.setdefault(stage, dict(dram_bytes=0.0, work_executed=0.0, warp_inst=0.0, time_us=0.0, launches=0, kernels=[]))
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            a["dram_bytes"] += val(r, m) * scale.get(units[col[m]], 1.0)
        cyc = val(r, "smsp__cycles_elapsed.max") or val(r, "sm__cycles_elapsed.max")
        rate = lambda op: val(r, "smsp__sass_thread_inst_executed_op_%s_pred_on.sum.per_cycle_elapsed" % op)
        a["work_executed"] += (2.0 * rate("ffma") + rate("fmul") + rate("fadd")) * cyc
        a["warp_inst"] += val(r, "smsp__inst_executed.sum")
        t = val(r, "gpu__time_duration.sum")
        a["time_us"] += t * {"us": 1.0, "ms": 1e3, "ns": 1e-3}.get(units[col["gpu__time_duration.sum"]], 1.0)
        a["launches"] += 1
        if name.split("(")[0] not in a["kernels"]:
            a["kernels"].append(name.split("(")[0][:60])
    for a in agg.values():
        for k in ("dram_bytes", "work_executed", "warp_inst"):
            a[k] = int(a[k])
        a["time_us"] = round(a["time_us"], 2)
    agg["_source"] = "ncu --set full --clock-control none of tests/gpu_profile_step.py (one fused mapping iteration, headline " \
                     "shape, cold caches per launch): " + path
    print(json.dumps(agg, indent=1))


if __name__ == "__main__":
    main(sys.argv[1])
