set -x
python bench.py --workload config4 --no-cpu-baseline > gpurun_out/r1g_c4_n1.log 2>&1; tail -c 1500 gpurun_out/r1g_c4_n1.log
python bench.py --workload config1 --no-cpu-baseline > gpurun_out/r1g_c1_n1.log 2>&1; tail -c 1500 gpurun_out/r1g_c1_n1.log
python bench.py --no-cpu-baseline > gpurun_out/r1g_c2_n1.log 2>&1; tail -c 1500 gpurun_out/r1g_c2_n1.log
