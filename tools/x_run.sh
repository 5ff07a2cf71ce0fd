timeout -s KILL 200 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "timemix" 2>&1 | tail -n 3 | cut -c1-300
timeout -s KILL 100 python tools/bench_kernels.py tm 2>&1 | tail -n 6
