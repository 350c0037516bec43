python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for T in q4_0 q8_0 f16; do python bench.py --kv-type $T --no-cpu-baseline --steps 30 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['dtype'], d['value'], d['roofline']['frac'], d['clocks'], d['e2e']['value'])"; done
python bench.py --workload config1 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('config1', d['value'], d['roofline']['frac'], d['ms_per_step'], d['e2e'])"
