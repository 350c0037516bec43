for T in q8_0; do python bench.py --kv-type $T --no-cpu-baseline --steps 30 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['dtype'], d['value'], d['roofline']['frac'], d['clocks'], d['e2e'])"; done
python bench.py --workload config1 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('config1', d['value'], d['roofline']['frac'], d['ms_per_step'], d['e2e'])"
