timeout 180 python -m pytest tests/test_gpu_prefill.py -x -q 2>&1 | tail -25
echo "rc=$?"
nvidia-smi --query-gpu=name,memory.used --format=csv,noheader
