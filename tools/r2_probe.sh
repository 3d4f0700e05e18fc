cd /root/repo
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_havac.py -x -q -m gpu 2>&1 | tail -3
for rows in 72 200; do HAVAC_KERNEL=f16x2 timeout 120 python tools/quick_bench.py 1e9 $rows 3 2>&1 | tail -1; done
timeout 100 python tools/quick_bench.py 1e8 35804 3 2>&1 | tail -1
