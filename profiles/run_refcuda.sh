#!/bin/bash
# profiles/run_refcuda.sh TAG — under gpurun: the reference's own operator test (tests/test-backend-ops.cpp, unmodified) against
# (a) the reference's CUDA backend built for sm_100 and (b) the same backend with FLASH_ATTN_EXT / CPY routed to
# libb200attn.so (oracle/Makefile target refcuda), then both timed on the dense-mask form of configs 1 and 2.
TAG=${1:-rX}
O=gpurun_out
R=oracle/_ref
mkdir -p $O
CTX2=$(python -c "import numpy as np; print(','.join(map(str, np.random.default_rng(7).integers(512, 8193, size=64))))")
CTX1=$(python -c "print(','.join(['1024']*8))")
for V in ref b200; do
  for OP in FLASH_ATTN_EXT CPY DUP CONT; do
    B200_GLUE_STATS=1 timeout 900 $R/test-backend-ops-$V test -o $OP -b CUDA0 > $O/${TAG}_tbo_${V}_${OP}.log 2>&1
    echo "$V $OP rc=$? $(grep -c 'OK' $O/${TAG}_tbo_${V}_${OP}.log) OK, $(grep -c 'FAIL' $O/${TAG}_tbo_${V}_${OP}.log) FAIL, $(grep -c 'not supported' $O/${TAG}_tbo_${V}_${OP}.log) not supported; $(grep 'b200 glue' $O/${TAG}_tbo_${V}_${OP}.log)"
    tail -3 $O/${TAG}_tbo_${V}_${OP}.log | head -2
  done
done
: > $O/${TAG}_refcuda_bench.jsonl
for V in ref b200; do
  for T in f16 q8_0 q4_0; do
    timeout 600 $R/ref_cuda_bench-$V --ctx $CTX1 --type $T --iters 200 2>> $O/${TAG}_refcuda_bench.err | sed "s/^{/{\"variant\": \"$V\", \"workload\": \"config1 8x1024\", /" >> $O/${TAG}_refcuda_bench.jsonl
    timeout 600 $R/ref_cuda_bench-$V --ctx $CTX2 --type $T --iters 30  2>> $O/${TAG}_refcuda_bench.err | sed "s/^{/{\"variant\": \"$V\", \"workload\": \"config2 64xU{512..8192}\", /" >> $O/${TAG}_refcuda_bench.jsonl
    timeout 300 $R/ref_cuda_bench-$V --cpy 512 --type $T --iters 200 2>> $O/${TAG}_refcuda_bench.err | sed "s/^{/{\"variant\": \"$V\", \"workload\": \"config3 KV store 512 tokens\", /" >> $O/${TAG}_refcuda_bench.jsonl
  done
done
cat $O/${TAG}_refcuda_bench.jsonl
tail -5 $O/${TAG}_refcuda_bench.err
