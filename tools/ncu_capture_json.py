#!/usr/bin/env python
"""profiles/ncu_capture.json from an `ncu --set full` report of bench.py (tools/gpu_ncu_capture.sh):
per kernel of the step, DRAM bytes and warp instructions of ONE launch, the worlds / agents of that launch, and a hash
of the kernel sources the report was taken from. bench.py reads the file for roofline.traffic / issue_frac and refuses
it (traffic: null, "STALE capture") when the sources have changed since.

  python tools/ncu_capture_json.py report.ncu-rep worlds agents [out.json]
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rep, worlds, agents = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    out_path = sys.argv[4] if len(sys.argv) > 4 else os.path.join(ROOT, "profiles", "ncu_capture.json")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h = rows[0]
    col = {n: i for i, n in enumerate(h)}
    units = rows[1]

    def val(r, name):
        v = float(r[col[name]].replace(",", ""))
        u = units[col[name]]
        return v * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0, "us": 1e-3, "ms": 1.0, "ns": 1e-6}.get(u, 1.0)

    kernels = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        fast = any(t in name for t in ("export_kernel<1>", "export_kernel<(bool)1>", "export_kernel<1,", "export_kernel<(bool)1,"))
        key = ("export_kernel" if fast else
               "step_kernel" if "step_kernel" in name else "shift_kernel" if "shift_kernel" in name else
               "scan_kernel" if "scan_kernel" in name else "tile_sum_kernel" if "tile_sum" in name else
               "action_write_kernel" if "action_write" in name else None)
        if key is None or key in kernels:
            continue
        kernels[key] = {
            "name": name, "worlds": worlds, "agents": agents,
            "duration_ms": val(r, "gpu__time_duration.sum"),
            "dram_read_bytes": val(r, "dram__bytes_read.sum"), "dram_write_bytes": val(r, "dram__bytes_write.sum"),
            "warp_instructions": val(r, "smsp__inst_executed.sum"),
            "threads_per_instruction": val(r, "smsp__thread_inst_executed_per_inst_executed.ratio"),
            "issue_active_pct": val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
            "registers": val(r, "launch__registers_per_thread"),
        }
    import bench
    doc = {"source_sha": bench.source_sha(), "report": os.path.basename(rep),
           "how": "ncu --set full --clock-control none, one launch per kernel of `bench.py --steps 5 --warmup 3` (cold caches, serialised: "
                  "durations are NOT bench values; bytes and instruction counts are per launch)",
           "kernels": kernels}
    with open(out_path, "w") as f:
        json.dump(doc, f, indent=1)
    print(json.dumps({k: {"ms": round(v["duration_ms"], 4), "MB": round((v["dram_read_bytes"] + v["dram_write_bytes"]) / 1e6, 1),
                          "Minst": round(v["warp_instructions"] / 1e6, 2)} for k, v in kernels.items()}))


if __name__ == "__main__":
    main()
