# Multi-GPU round on N GPUs of one box: frame-sharded bench line, then the z-sharded dense grid with the NCCL
# combine and with the peer-memory combine (bit-identity to the unsharded planes is checked inside).
#   gpurun --gpus N -- bash tools/multi_gpu_round.sh N TAG
N=${1:-2}
TAG=${2:-mg}
mkdir -p gpurun_out
launch() {  # launch PORT script args...
  local port=$1; shift
  if [ "$N" = "1" ]; then python "$@"; else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port "$@"; fi
}
timeout 600 bash -c "$(declare -f launch); N=$N; launch 29501 bench.py --gpus $N --steps 20 --warmup 3" > gpurun_out/${TAG}_bench_${N}gpu.json 2> gpurun_out/${TAG}_bench_${N}gpu.err
PORT=29510
for P in 0 pull multimem; do
  PORT=$((PORT + 1))
  PEER=$P timeout 300 bash -c "$(declare -f launch); N=$N; launch $PORT tools/zslab_check.py" 2> gpurun_out/${TAG}_zslab_${N}gpu_peer$P.err | grep '^{' > gpurun_out/${TAG}_zslab_${N}gpu_peer$P.json
done
cut -c1-200 gpurun_out/${TAG}_bench_${N}gpu.json; cat gpurun_out/${TAG}_zslab_${N}gpu_peer*.json
