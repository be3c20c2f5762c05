python -m pytest tests -m gpu -x -q > gpurun_out/r15_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r15_pytest.log
python bench.py --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/r15_bench.json 2>gpurun_out/r15_bench.err
python -c "
import json; d=json.load(open('gpurun_out/r15_bench.json')); print('bench', d['value'], d['e2e']['value'], {k:v['ms'] for k,v in d['stages'].items()})"
ncu --metrics gpu__time_duration.sum,sm__inst_executed.sum --clock-control none --profile-from-start off -k regex:"onesweep|duplicate|ranges" --csv --log-file gpurun_out/r15_frame.csv python tools/prof_frame.py human_scene_1080p 0 > gpurun_out/r15.log 2>&1
