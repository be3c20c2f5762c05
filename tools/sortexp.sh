python -m pytest tests -m gpu -x -q > gpurun_out/r21_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r21_pytest.log
python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r21_bench.json 2>gpurun_out/r21_bench.err
python -c "
import json; d=json.load(open('gpurun_out/r21_bench.json')); print('bench', d['value'], d['e2e']['value'], {k:v['ms'] for k,v in d['stages'].items()})"
