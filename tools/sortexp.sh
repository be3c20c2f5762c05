python -m pytest tests -m gpu -x -q > gpurun_out/r7_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r7_pytest.log
for sh in 0 1 2 3; do
  HB_SORT_SHAPE=$sh python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r7_shape$sh.json 2>gpurun_out/r7_shape$sh.err
  python -c "
import json; d=json.load(open('gpurun_out/r7_shape$sh.json')); print('shape $sh', d['value'], {k:v['ms'] for k,v in d['stages'].items()})"
done
for sh in 0 1; do
  HB_SORT_NOLOOK=1 HB_SORT_SHAPE=$sh python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r7_nolook$sh.json 2>gpurun_out/r7_nolook$sh.err
  python -c "
import json; d=json.load(open('gpurun_out/r7_nolook$sh.json')); print('nolook shape $sh', d['value'], 'sort', d['stages']['sort_pairs']['ms'])"
done
