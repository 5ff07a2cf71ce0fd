timeout -s KILL 300 python bench.py --no-cpu --no-gpu-baseline > gpurun_out/r2q_bench.json 2>/dev/null; python -c "
import json;d=json.loads(open('gpurun_out/r2q_bench.json').read().strip().splitlines()[-1]);print(d['value'],d['ms_per_step'],d['e2e'],d['roofline']['frac'],d['clocks']);[print(k,v['ms_per_step'],v['launches_per_step'],v['tflops'],v['gbs']) for k,v in d['kernels'].items()]"
