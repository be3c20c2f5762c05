#!/bin/bash
TAG=${1:-rXX}; O=gpurun_out; mkdir -p $O
timeout 1700 python -m pytest tests -x -q -m gpu > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu-baseline > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('$O/${TAG}_bench.json')); print('  value %.2f e2e %.2f' % (d['value'], d['e2e']['value']), {k: v['ms'] for k, v in d['stages'].items()})"
