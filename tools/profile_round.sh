# Round profile (one GPU): clean bench, then the ncu launch list of the timed region of the same command, then one
# full capture of the per-person projection kernel (project_walk3_kernel, the launch with 320 cubes) and one of the
# tcgen05 residual block (res2d_block_kernel, the launch with 960 planes) and of the tcgen05 7x7 front convolution.
#   bash tools/profile_round.sh TAG [full]
set -x
TAG=${1:-prof}
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || exit 1
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-extras > gpurun_out/${TAG}_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-extras > gpurun_out/${TAG}_ncu.log 2>&1
if [ "$2" = "full" ]; then
# walk launches of a step: the space grid, the z columns of the top-k cells (cube form, tiny), the person cubes: -s 2 picks the last
ncu --set full --clock-control none --import-source on -k regex:project_walk3_kernel --profile-from-start off -s 2 -c 1 -o gpurun_out/${TAG}_full_walk3 python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-extras > gpurun_out/${TAG}_ncu_full.log 2>&1
# the tcgen05 residual block: CenterNet's launch comes first in a step, P2PNet's (960 planes) second
ncu --set full --clock-control none --import-source on -k regex:res2d_block_kernel --profile-from-start off -s 1 -c 1 -o gpurun_out/${TAG}_full_res2d python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-extras > gpurun_out/${TAG}_ncu_full_res2d.log 2>&1
# the tcgen05 7x7 front convolution: CenterNet's launch first, P2PNet's (960 planes) second
ncu --set full --clock-control none --import-source on -k regex:conv7x7_tc_kernel --profile-from-start off -s 1 -c 1 -o gpurun_out/${TAG}_full_conv7tc python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-extras > gpurun_out/${TAG}_ncu_full_conv7tc.log 2>&1
# the tcgen05 tail kernel of P2PNet (one launch per step)
ncu --set full --clock-control none --import-source on -k regex:res_tail_kernel --profile-from-start off -s 0 -c 1 -o gpurun_out/${TAG}_full_restail python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-extras > gpurun_out/${TAG}_ncu_full_restail.log 2>&1
fi
ls -la gpurun_out | tail -6
