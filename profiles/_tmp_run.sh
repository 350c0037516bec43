timeout 400 python -m pytest tests/test_gpu_prefill.py -x -q 2>&1 | tail -6
for T in q8_0 q4_0; do timeout 300 python bench.py --workload config3 --kv-type $T --layers 8 --steps 10 2>&1 | tail -1 | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['attention_only'], d['kv_write_only'])"; done
