timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3; timeout 300 python bench.py --no-cpu-baseline > gpurun_out/bench2.json 2> gpurun_out/bench2.err; python -c "
import json;d=json.load(open('gpurun_out/bench2.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d['gpu_launches']);print({k:v['ms'] for k,v in d['kernels'].items()})"; tail -3 gpurun_out/bench2.err
