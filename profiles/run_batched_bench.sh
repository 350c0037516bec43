#!/bin/bash
# profiles/run_batched_bench.sh TAG [NPL] — under gpurun: BASELINE config 5 / SURVEY 8(f) N1 on ONE GPU.  The reference's
# own llama-batched-bench (examples/batched-bench, unmodified) on a random-init Llama-3-8B Q4_0 GGUF with a Q4_0 KV cache
# and flash attention, three ways:
#   ref     the reference's CUDA backend, every node on its own kernels                      (B200_GLUE=0)
#   b200    same libllama, same graphs (dense mask); FLASH_ATTN_EXT / CPY on libb200attn.so  (B200_GLUE=1)
#   paged   libllama with the paged KV section (llama.cpp_test_b200/host/libllama_paged.patch): cells from PagedKVCache,
#           ONE paged FLASH_ATTN_EXT node per layer with the fused slot-mapped KV store       (LLAMA_KV_PAGED=1)
TAG=${1:-rX}
NPL=${2:-256}
O=gpurun_out
R=oracle/_ref
mkdir -p $O
LAYERS=${LAYERS:-32}
M=/tmp/llama3_8b_q4_0_random_${LAYERS}l.gguf
[ -f $M ] || python profiles/make_random_gguf.py $M --layers $LAYERS > $O/${TAG}_gguf.log 2>&1
ls -la $M
run() {  # name, binary, env..., then -- args
  local name=$1 bin=$2; shift 2
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" GGML_NO_BACKTRACE=1 B200_GLUE_STATS=1 B200_GLUE_TIMING=1 timeout -k 5 ${RUN_TIMEOUT:-300} $R/$bin -m $M -ngl 99 -fa -ctk q4_0 -ctv q4_0 -b 2048 -ub 512 "$@" > $O/${TAG}_bb_${name}.log 2>&1
  echo "== $name rc=$?"; grep -E "^\|" $O/${TAG}_bb_${name}.log | tail -4; grep "b200 glue" $O/${TAG}_bb_${name}.log
}
for SHAPE in ${SHAPES:-"128 128" "896 128"}; do
  set -- $SHAPE; PP=$1; TG=$2; CTX=$(( NPL*(PP+TG) ))
  S=pp${PP}_tg${TG}_npl${NPL}
  case " ${MODES:-ref b200 paged} " in *" ref "*)   run ref_$S   llama-batched-bench       B200_GLUE=0 -- -c $CTX -npp $PP -ntg $TG -npl $NPL;; esac
  case " ${MODES:-ref b200 paged} " in *" b200 "*)  run b200_$S  llama-batched-bench       B200_GLUE=1 -- -c $CTX -npp $PP -ntg $TG -npl $NPL;; esac
  case " ${MODES:-ref b200 paged} " in *" paged "*) run paged_$S llama-batched-bench-paged B200_GLUE=1 LLAMA_KV_PAGED=1 -- -c $CTX -npp $PP -ntg $TG -npl $NPL;; esac
done
