#!/bin/bash
# profiles/run_scale.sh TAG N — multi-GPU bench lines on one box (run under gpurun --gpus N): config 2 sequence-sharded
# (weak scaling, no collective) and config 4 KV-head sharded (strong scaling, NCCL all-gather per layer).
TAG=${1:-rX}; N=${2:-8}
O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -8 > $O/${TAG}_n${N}_gpus.txt
for WL in config2 config4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --workload $WL --steps 30 --warmup 5 --no-cpu-baseline > $O/${TAG}_${WL}_n${N}.log 2>&1
  tail -1 $O/${TAG}_${WL}_n${N}.log | cut -c1-900; echo
done
