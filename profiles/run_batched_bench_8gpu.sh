#!/bin/bash
# profiles/run_batched_bench_8gpu.sh TAG — under `gpurun --gpus 8`: BASELINE config 5 as stated — llama-batched-bench, random-init
# Llama-3-8B Q4_0, 256 parallel sequences, sequence-batch sharded over 8 x B200 = eight independent replicas of 32 sequences
# (SURVEY 8e partition A: no collective), all running at once; the aggregate S_TG / S_PP is the sum over the replicas.
# Modes as in run_batched_bench.sh: ref (reference CUDA kernels), paged (libllama_paged.patch + libb200attn).
TAG=${1:-rX}
NG=${NG:-8}
O=gpurun_out; R=oracle/_ref; mkdir -p $O
M=/tmp/llama3_8b_q4_0_random_32l.gguf
[ -f $M ] || python profiles/make_random_gguf.py $M --layers 32 > $O/${TAG}_gguf.log 2>&1
ls -la $M
PP=896; TG=128; NPL=32; CTX=$(( NPL*(PP+TG) ))
for MODE in ref paged; do
  if [ $MODE = ref ]; then BIN=llama-batched-bench; ENVS="B200_GLUE=0"; else BIN=llama-batched-bench-paged; ENVS="B200_GLUE=1 LLAMA_KV_PAGED=1"; fi
  pids=()
  for g in $(seq 0 $((NG-1))); do
    env $ENVS CUDA_VISIBLE_DEVICES=$g GGML_NO_BACKTRACE=1 timeout -k 5 400 $R/$BIN -m $M -ngl 99 -fa -ctk q4_0 -ctv q4_0 -b 2048 -ub 512 \
        -c $CTX -npp $PP -ntg $TG -npl $NPL > $O/${TAG}_bb8_${MODE}_gpu$g.log 2>&1 &
    pids+=($!)
  done
  for p in "${pids[@]}"; do wait $p; done
  python - "$O" "$TAG" "$MODE" "$NG" <<'PY' | tee -a $O/${TAG}_bb8_summary.txt
import sys, re
o, tag, mode, ng = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4])
spp = stg = 0.0; rows = []
for g in range(ng):
    for line in open(f"{o}/{tag}_bb8_{mode}_gpu{g}.log", errors="replace"):
        m = re.match(r"\|\s*(\d+)\s*\|\s*(\d+)\s*\|\s*(\d+)\s*\|\s*(\d+)\s*\|\s*([\d.]+)\s*\|\s*([\d.]+)\s*\|\s*([\d.]+)\s*\|\s*([\d.]+)\s*\|", line)
        if m:
            spp += float(m.group(6)); stg += float(m.group(8)); rows.append((g, float(m.group(6)), float(m.group(8))))
print(f"{mode}: {len(rows)} of {ng} replicas reported; aggregate S_PP {spp:.1f} t/s, S_TG {stg:.1f} t/s; per GPU S_TG " + " ".join(f"{r[2]:.0f}" for r in rows))
PY
done
