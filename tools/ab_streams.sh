for S in 1 2 3; do HB_STREAMS=$S python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/ab_$S.json 2> gpurun_out/ab_$S.err; tail -2 gpurun_out/ab_$S.err; python -c "
import json; d=json.load(open('gpurun_out/ab_$S.json')); print('streams $S value', d['value'], 'e2e', d['e2e']['value'])"; done
