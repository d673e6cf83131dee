#!/usr/bin/env python
"""profiles/ncu_constants.json from an `ncu --set full` capture of `bench.py --profile`:

  ncu -i gpurun_out/<rep>.ncu-rep --page raw --csv > profiles/<name>_full_raw.csv
  python scripts/ncu_constants.py profiles/<name>_full_raw.csv <git sha the capture was built from>

bench.py reads the file for the two figures it cannot measure live: warp instructions and DRAM bytes per launch of
the two fused kernels at the C2 workload (512 x 600 x 1000).  The SHA says which build they belong to."""
import csv
import json
import os
import sys

KERNEL_OF = {"k_fused3<3, 2, 4, 0>": "sap_mean_var0.4", "k_fused3<1, 1, 4, 0>": "gaussian_gaus_blur_var1.5"}


def main():
    src, sha = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(open(src)))
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}
    out = {"source": os.path.relpath(src), "git_sha": sha, "workload": "C2: 512x600x1000", "kernels": {}}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        spec = next((s for k, s in KERNEL_OF.items() if k in name), None)
        if spec is None:
            continue
        out["kernels"][spec] = {
            "kernel": name.split("(")[0].replace("void ", ""),
            "warp_instructions": float(r[col["smsp__inst_executed.sum"]]),
            "dram_bytes": (float(r[col["dram__bytes_read.sum"]]) + float(r[col["dram__bytes_write.sum"]])) * 1e6,
            "ncu_us": float(r[col["gpu__time_duration.sum"]]),
            "issue_active_pct": float(r[col["smsp__issue_active.avg.pct_of_peak_sustained_active"]]),
            "alu_pct": float(r[col["sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"]]),
            "fma_pct": float(r[col["sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"]]),
            "xu_pct": float(r[col["sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"]]),
        }
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_constants.json")
    json.dump(out, open(path, "w"), indent=1, sort_keys=True)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
