python -m pytest tests/test_gpu_parity.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3
for T in q4_0 q8_0 f16; do python bench.py --kv-type $T --no-cpu-baseline --steps 30 --no-e2e 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['dtype'], d['value'], d['roofline']['frac'], d['clocks'])"; done
