timeout -s KILL 300 python bench.py --no-cpu --no-gpu-baseline > gpurun_out/r2r_bench.json 2>gpurun_out/r2r_bench.err; tail -n 3 gpurun_out/r2r_bench.err; python -c "
import json;d=json.loads(open('gpurun_out/r2r_bench.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['e2e'],d['roofline']['frac'])"
