python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for T in q4_0 q8_0; do python bench.py --kv-type $T --no-cpu-baseline --steps 30 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['dtype'], d['value'], d['roofline']['frac'], d['clocks'], d['e2e']['value'])"; done
ncu --set full --clock-control none --import-source on -k regex:sk_main -s 4 -c 1 -o gpurun_out/r1i_sk_q4_0 -f python bench.py --kv-type q4_0 --layers 2 --steps 2 --warmup 3 --no-graph --no-cpu-baseline --no-e2e > gpurun_out/r1i_ncu_q4_0.log 2>&1
