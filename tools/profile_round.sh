set -x
python bench.py --steps 20 --warmup 3 > gpurun_out/r21_bench.json 2> gpurun_out/r21_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r21_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r21_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:project_fused_kernel --profile-from-start off -c 1 -o gpurun_out/r21_full_fused python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r21_ncu_full.log 2>&1
ls -la gpurun_out | tail
