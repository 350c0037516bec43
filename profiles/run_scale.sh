#!/bin/bash
# profiles/run_scale.sh TAG N — multi-GPU bench lines on one box (run under gpurun --gpus N): config 2 sequence-sharded
# (weak scaling, no collective) and config 4 KV-head sharded (strong scaling, one output exchange per layer: the
# library's own NVLink peer-memory kernels, and NCCL all-gather for comparison).
TAG=${1:-rX}; N=${2:-8}
O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -8 > $O/${TAG}_n${N}_gpus.txt
run() {  # name, bench args...
  local name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline "$@" > $O/${TAG}_${name}_n${N}.log 2>&1
  tail -1 $O/${TAG}_${name}_n${N}.log | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('$name', 'N', d['n_gpus'], d['value'], d['unit'], 'ms', d['ms_per_step'], 'frac', d['roofline']['frac'], 'e2e', d.get('e2e',{}).get('value'), d['config'].get('exchange','')[:40])" || tail -5 $O/${TAG}_${name}_n${N}.log
}
run config4_peer --workload config4 --exchange peer
run config4_nccl --workload config4 --exchange nccl --no-e2e
run config2 --workload config2
