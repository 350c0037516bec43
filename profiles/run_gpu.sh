#!/bin/bash
# profiles/run_gpu.sh TAG — the round's standard GPU pass (run under gpurun): parity tests, bench lines for the
# three KV types + the reference arm, the ncu launch list and one --set full capture of the dominant kernel.
TAG=${1:-rX}
O=gpurun_out
mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" >> $O/${TAG}_pytest.log
tail -3 $O/${TAG}_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; tail -1 $O/${TAG}_smoke.log
for T in q8_0 f16 q4_0; do
  python bench.py --kv-type $T > $O/${TAG}_bench_$T.log 2>&1; tail -c 600 $O/${TAG}_bench_$T.log; echo
done
for WL in config1 config3 config4 config5_attn; do
  python bench.py --workload $WL --no-cpu-baseline > $O/${TAG}_bench_$WL.log 2>&1; tail -c 400 $O/${TAG}_bench_$WL.log; echo
done
for T in f16 q8_0 q4_0; do   # prefill, un-capped short runs per cache type (quantised caches go through the F16 scratch)
  python bench.py --workload config3 --kv-type $T --layers 8 --steps 10 --no-cpu-baseline > $O/${TAG}_bench_config3_${T}_8l.log 2>&1; tail -c 300 $O/${TAG}_bench_config3_${T}_8l.log; echo
done
python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_reference.log 2>&1; tail -c 400 $O/${TAG}_bench_reference.log; echo
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --layers 4 --steps 2 --warmup 3 --no-graph --no-cpu-baseline > $O/${TAG}_ncu_list.log 2>&1
for T in q8_0 q4_0 f16; do
  ncu --set full --clock-control none --import-source on -k regex:sk_main -s 4 -c 1 -o $O/${TAG}_sk_$T -f \
    python bench.py --kv-type $T --layers 2 --steps 2 --warmup 3 --no-graph --no-cpu-baseline > $O/${TAG}_ncu_$T.log 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:pf_main -s 2 -c 1 -o $O/${TAG}_pf -f \
    python bench.py --workload config3 --layers 2 --steps 2 --no-graph > $O/${TAG}_ncu_pf.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/${TAG}_launches_config3.csv \
    python bench.py --workload config3 --layers 2 --steps 2 --no-graph > $O/${TAG}_ncu_list3.log 2>&1
for WL in config1 config4_shard8; do   # the latency-bound ops: where the ~10 us of launch chain around 5 / 80 us of bytes go
  ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/${TAG}_launches_$WL.csv \
      python bench.py --workload $WL --layers 4 --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-e2e > $O/${TAG}_ncu_list_$WL.log 2>&1
done
python profiles/bench_prefill_single.py > $O/${TAG}_prefill_single.jsonl 2>&1; cat $O/${TAG}_prefill_single.jsonl   # one sequence's chunk: the KV split of the tail
bash profiles/run_timeline.sh $TAG config1 config5_attn > $O/${TAG}_timeline.log 2>&1; grep -A3 "^==" $O/${TAG}_timeline.log   # needs gpurun_variants/tl (make timeline)
ls -la $O | tail -15
