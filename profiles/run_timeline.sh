#!/bin/bash
# profiles/run_timeline.sh TAG [WORKLOAD...] — under gpurun: where the time of a small layer-op goes.  Runs bench.py's
# graph-replayed decode loop on the -DB200_TIMELINE build (make -C llama.cpp_test_b200/csrc timeline, done in the
# container) and prints the steady-state timeline of one layer-op (profiles/timeline_report.py).  Not a bench number.
TAG=${1:-rX}; shift
O=gpurun_out; mkdir -p $O
for WL in ${@:-config1 config5_attn}; do
  B200_LIB=$PWD/gpurun_variants/tl/libb200attn.so B200_TIMELINE_OUT=$O/${TAG}_timeline_$WL.txt \
    timeout 300 python bench.py --workload $WL --steps 20 --warmup 3 --no-cpu-baseline --no-e2e > $O/${TAG}_timeline_$WL.log 2>&1
  echo "== $WL"; python profiles/timeline_report.py $O/${TAG}_timeline_$WL.txt | tee $O/${TAG}_timeline_${WL}_report.txt
done
