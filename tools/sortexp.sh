python bench.py --steps 30 --warmup 3 --no-cpu-baseline > gpurun_out/r23_bench.json 2>gpurun_out/r23_bench.err; tail -3 gpurun_out/r23_bench.err
python -c "
import json; d=json.load(open('gpurun_out/r23_bench.json')); print('bench', d['value'], d['e2e']['value'], d['ms_per_step'], d['frame'], {k:v['ms'] for k,v in d['stages'].items()})"
python bench.py --impl reference --steps 5 --warmup 3 | cut -c1-200
