"""Summarise `ncu --set full` reports into profiles/: one row per captured kernel.
    python tools/ncu_summary.py <tag> <version>      # reads gpurun_out/<tag>_full_*.ncu-rep
Writes profiles/r01_ncu_kernels_<version>.csv, the full metric list of the per-person projection kernel
(profiles/r01_ncu_full_project_fused_<version>.csv) and refreshes profiles/r01_ncu_summary.json."""
import csv
import glob
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0, "us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}
KEYS = [("gpu__time_duration.sum", "duration_ms"), ("dram__bytes_read.sum", "dram_read_bytes"),
        ("dram__bytes_write.sum", "dram_write_bytes"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1_pct"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2_pct"),
        ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue_pct"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy_pct"),
        ("l1tex__t_sector_hit_rate.pct", "l1_hit_pct"), ("lts__t_sector_hit_rate.pct", "l2_hit_pct"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "warp_insts")]


def raw(path):
    txt = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    return rows[0], rows[1], rows[2]


def num(v, u):
    return float(v.replace(",", "")) * UNIT.get(u, 1.0)


def main(tag, version):
    out_rows = []
    for path in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", f"{tag}_full_*.ncu-rep"))):
        hdr, units, vals = raw(path)
        d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
        row = {"report": os.path.basename(path), "kernel": d["Kernel Name"][0][:70], "grid": d["Grid Size"][0],
               "block": d["Block Size"][0]}
        for k, name in KEYS:
            row[name] = num(*d[k]) if k in d and d[k][0] not in ("", "n/a") else None
        out_rows.append(row)
        if path.endswith(f"{tag}_full_fused.ncu-rep"):
            with open(os.path.join(ROOT, "profiles", f"r01_ncu_full_project_fused_{version}.csv"), "w") as f:
                w = csv.writer(f)
                w.writerow(["metric", "unit", "value"])
                for h, u, v in zip(hdr, units, vals):
                    w.writerow([h, u, v])
            summ = {"kernel": "project_fused_kernel<4,0,1> grid (1024,320,1): 320 person cubes of 64^3 -> 3 max-planes, "
                              "Panoptic, B=32, heatmaps G, bbox full",
                    "source": f"profiles/r01_ncu_full_project_fused_{version}.csv (ncu --set full --clock-control none, "
                              "tools/profile_round.sh)",
                    "dram_bytes_read": row["dram_read_bytes"], "dram_bytes_write": row["dram_write_bytes"],
                    "traffic_bytes_per_launch": row["dram_read_bytes"] + row["dram_write_bytes"],
                    "issue_active_pct": row["issue_pct"], "l1_hit_pct": row["l1_hit_pct"], "l2_hit_pct": row["l2_hit_pct"],
                    "warps_active_pct": row["occupancy_pct"], "registers_per_thread": row["regs"],
                    "duration_ms": row["duration_ms"]}
            with open(os.path.join(ROOT, "profiles", "r01_ncu_summary.json"), "w") as f:
                json.dump(summ, f, indent=1)
    cols = ["report", "kernel", "grid", "block"] + [n for _, n in KEYS]
    with open(os.path.join(ROOT, "profiles", f"r01_ncu_kernels_{version}.csv"), "w") as f:
        w = csv.DictWriter(f, fieldnames=cols)
        w.writeheader()
        for r in out_rows:
            w.writerow(r)
    for r in out_rows:
        tr = (r["dram_read_bytes"] or 0) + (r["dram_write_bytes"] or 0)
        f = lambda v: "n/a" if v is None else f"{v:.0f}"
        print(f"{r['kernel'][:44]:44s} {r['grid']:>16s} {r['duration_ms']:8.3f} ms  dram {tr / 1e6:8.1f} MB "
              f"({tr / max(r['duration_ms'], 1e-9) / 1e6:7.1f} GB/s, {f(r['dram_pct'])}%)  sm {f(r['sm_pct'])}%  "
              f"l1 {f(r['l1_pct'])}%  issue {f(r['issue_pct'])}%  occ {f(r['occupancy_pct'])}%  regs {f(r['regs'])}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "v5")
