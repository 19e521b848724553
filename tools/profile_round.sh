# Round profile: clean bench, then the ncu launch list of the timed region of the same command, then one
# full capture of the per-person projection kernel (third project_fused launch of a step: planes, columns, persons).
set -x
TAG=${1:-prof}
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 3 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${TAG}_launches.csv python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_ncu.log 2>&1
if [ "$2" = "full" ] || [ "$2" = "fused" ]; then
ncu --set full --clock-control none --import-source on -k regex:project_fused_kernel --profile-from-start off -s 2 -c 1 -o gpurun_out/${TAG}_full_fused python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_ncu_full.log 2>&1
fi
if [ "$2" = "full" ]; then
# the other hand-written kernels of the step, one full capture each (first launch inside the timed region)
for K in rowpair_kernel weightnet_features softargmax nms_topk_kernel pointwise_planar_kernel maxpool2_nhwc_kernel upshuffle_add_kernel; do
ncu --set full --clock-control none --import-source on -k regex:$K --profile-from-start off -c 1 -o gpurun_out/${TAG}_full_$K python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_ncu_$K.log 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:project_fused_kernel --profile-from-start off -c 1 -o gpurun_out/${TAG}_full_fused_space python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_ncu_fused_space.log 2>&1
fi
ls -la gpurun_out | tail -5
