#!/bin/bash
# what the driver runs at round end, in its order: build + smoke, the GPU tests, the default bench line, the reference arm
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.build(); g.smoke(); print('smoke ok')" > gpurun_out/fin_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/fin_smoke.log
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/fin_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/fin_pytest.log
timeout 900 python bench.py > gpurun_out/fin_bench.json 2> gpurun_out/fin_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/fin_ref.json 2> gpurun_out/fin_ref.err; echo "ref rc=$?"
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/fin_bench.json') if l.startswith('{')][-1])
print(d['ms_per_step'], d['value'], d['e2e']['value'], d['gpu_launches'], d['clocks'])
r=json.loads([l for l in open('gpurun_out/fin_ref.json') if l.startswith('{')][-1])
print(r.get('impl'), r.get('value'), r.get('config',{}).get('workload','')[:60], r.get('cpu_baseline'))
PY
