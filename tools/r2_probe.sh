cd /root/repo
echo "== main rows 1024"; timeout 60 python tools/quick_bench.py 24576 1024 1 1 2>&1 | tail -1
echo "== main C2"; timeout 100 python tools/quick_bench.py 1e8 35804 3 2>&1 | tail -2
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_havac.py -x -q -m gpu 2>&1 | tail -3
