# GPU suite against a build with device-side bounds assertions (-DFVP_PARANOID): stands in for compute-sanitizer memcheck,
# which is closed on this pool.  Build the library first where nvcc lives:  python -m faster_voxelpose_b200.build --paranoid
#   bash tools/paranoid_check.sh  -> gpurun_out/paranoid_check.txt
mkdir -p gpurun_out
FVP_B200_LIB=$PWD/faster_voxelpose_b200/libfvp_b200_paranoid.so python -m pytest tests -m gpu -q -x \
  -k "not streamed and not graphed and not whole_model and not fused_inference" > gpurun_out/paranoid_check.txt 2>&1
echo "exit code $?" >> gpurun_out/paranoid_check.txt
grep -c "FVP_ASSERT failed" gpurun_out/paranoid_check.txt | sed 's/^/FVP_ASSERT failures: /' >> gpurun_out/paranoid_check.txt
tail -5 gpurun_out/paranoid_check.txt
