#!/bin/bash
# profiles/run_logits_check.sh TAG — under gpurun: the same libllama call sequence (oracle/llama_logits_dump.cpp: interleaved
# prompts, a fork, decode steps, a dropped and a truncated sequence) three ways, logits compared with the reference run:
#   ref    reference CUDA kernels (B200_GLUE=0)      b200   libb200attn on the unmodified dense-mask graphs
#   paged  libllama with the paged KV section (libllama_paged.patch, LLAMA_KV_PAGED=1)
TAG=${1:-rX}
O=gpurun_out; R=oracle/_ref; mkdir -p $O
M=/tmp/llama3_8b_q4_0_random_4l.gguf
[ -f $M ] || python profiles/make_random_gguf.py $M --layers 4 > $O/${TAG}_gguf4.log 2>&1
for T in q4_0 q8_0 f16; do
  GGML_NO_BACKTRACE=1 B200_GLUE=0 timeout -k 5 120 $R/llama-logits-dump $M /tmp/l_ref_$T.bin -ctk $T > $O/${TAG}_logits_ref_$T.log 2>&1
  GGML_NO_BACKTRACE=1 B200_GLUE=1 B200_GLUE_STATS=1 timeout -k 5 120 $R/llama-logits-dump $M /tmp/l_b200_$T.bin -ctk $T > $O/${TAG}_logits_b200_$T.log 2>&1
  GGML_NO_BACKTRACE=1 B200_GLUE=1 B200_GLUE_STATS=1 LLAMA_KV_PAGED=1 timeout -k 5 120 $R/llama-logits-dump-paged $M /tmp/l_paged_$T.bin -ctk $T > $O/${TAG}_logits_paged_$T.log 2>&1
  grep -h "b200 glue:" $O/${TAG}_logits_b200_$T.log $O/${TAG}_logits_paged_$T.log
done
python - <<PY | tee $O/${TAG}_logits_check.txt
import numpy as np
for t in ("q4_0", "q8_0", "f16"):
    ref = np.fromfile(f"/tmp/l_ref_{t}.bin", dtype=np.float32)
    for v in ("b200", "paged"):
        try:
            x = np.fromfile(f"/tmp/l_{v}_{t}.bin", dtype=np.float32)
        except Exception as e:
            print(t, v, "missing", e); continue
        if x.size != ref.size or ref.size == 0:
            print(t, v, "size mismatch", x.size, ref.size); continue
        rows = ref.size // 128256
        r, y = ref.reshape(rows, -1), x.reshape(rows, -1)
        nmse = ((r - y) ** 2).sum(1) / (r ** 2).sum(1)
        same_top = (r.argmax(1) == y.argmax(1)).mean()
        print(f"kv {t:5s} {v:5s} vs reference kernels: rows {rows}, logits NMSE max {nmse.max():.3e} mean {nmse.mean():.3e}, "
              f"same argmax in {same_top*100:.0f}% of rows, finite {np.isfinite(y).all()}")
PY
