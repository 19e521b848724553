# Copy one tools/profile_round.sh TAG full round from gpurun_out/ into profiles/ (bench line, launch list + summary, the raw
# metric lists and key-metric summaries of the four captured kernels) and refresh profiles/r02_ncu_summary.json.
#   bash tools/refresh_profiles.sh TAG
TAG=${1:?tag}
R=r02
cp gpurun_out/${TAG}_bench.json profiles/${R}_bench_1gpu.json
cp gpurun_out/${TAG}_launches.csv profiles/${R}_launches.csv
python tools/launch_summary.py gpurun_out/${TAG}_launches.csv | sed "s#gpurun_out/${TAG}_launches.csv#profiles/${R}_launches.csv#" > profiles/${R}_launches_summary.txt
for k in walk3:project_walk3:walk3 res2d:res2d_block:res2d conv7tc:conv7x7_tc:conv7x7_tc restail:res_tail:res_tail; do
  a=${k%%:*}; rest=${k#*:}; b=${rest%%:*}; c=${rest##*:}
  ncu -i gpurun_out/${TAG}_full_$a.ncu-rep --page raw --csv 2>/dev/null > profiles/${R}_ncu_full_$b.csv
  bash tools/ncu_key.sh gpurun_out/${TAG}_full_$a.ncu-rep > profiles/${R}_ncu_${c}_summary.txt
done
for d in shelf campus; do [ -f gpurun_out/${TAG}_bench_$d.json ] && cp gpurun_out/${TAG}_bench_$d.json profiles/${R}_bench_$d.json; done
python - "$TAG" <<'PY'
import json, re, sys
tag = sys.argv[1]
d = json.load(open('profiles/r02_ncu_summary.json'))
U = {"us": 1e-3, "ms": 1.0, "Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3}
def parse(path):
    out = {}
    for l in open(path):
        m = re.match(r"\s+(\S+)\s+([\d.,]+)\s*(\S*)", l)
        if m:
            out[m.group(1)] = (float(m.group(2).replace(',', '')), m.group(3))
    return out
g = lambda t, k: t[k][0] * U.get(t[k][1], 1.0)
w = parse('profiles/r02_ncu_walk3_summary.txt')
d["source"] = f"profiles/r02_ncu_full_project_walk3.csv (ncu --set full --clock-control none, tools/profile_round.sh {tag} full)"
d["dram_bytes_read"] = g(w, "dram__bytes_read.sum"); d["dram_bytes_write"] = g(w, "dram__bytes_write.sum")
d["traffic_bytes_per_launch"] = d["dram_bytes_read"] + d["dram_bytes_write"]
d["issue_active_pct"] = w["smsp__issue_active.avg.pct_of_peak_sustained_active"][0]
d["l1_hit_pct"] = w["l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct"][0]
d["l2_hit_pct"] = w["lts__t_sector_hit_rate.pct"][0]
d["warps_active_pct"] = w["sm__warps_active.avg.pct_of_peak_sustained_active"][0]
d["duration_ms"] = g(w, "gpu__time_duration.sum")
for key, path in (("res2d_block_kernel<16>", 'profiles/r02_ncu_res2d_summary.txt'), ("conv7x7_tc_kernel<16>", 'profiles/r02_ncu_conv7x7_tc_summary.txt'),
                  ("res_tail_kernel<16>", 'profiles/r02_ncu_res_tail_summary.txt')):
    r = parse(path)
    d[key].update({"duration_ms": g(r, "gpu__time_duration.sum"), "dram_bytes_read": g(r, "dram__bytes_read.sum"),
                   "dram_bytes_write": g(r, "dram__bytes_write.sum"),
                   "issue_active_pct": r["smsp__issue_active.avg.pct_of_peak_sustained_active"][0],
                   "warp_instructions": r["smsp__inst_executed.sum"][0]})
json.dump(d, open('profiles/r02_ncu_summary.json', 'w'), indent=1)
print("walk3 capture", d["duration_ms"], "ms")
PY
