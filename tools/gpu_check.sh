#!/bin/bash
# quick gate: the whole GPU suite, smoke, the default bench line
T=${1:-chk}; mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${T}_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/${T}_bench.log 2>&1; echo "bench rc=$?"
python - gpurun_out/${T}_bench.log <<'P'
import sys, json
d = json.loads([l for l in open(sys.argv[1]) if l.startswith('{')][-1])
print(round(d['value'], 1), d['config']['variant_resolved'], 'ms', round(d['ms_per_step'], 2), 'frac', round(d['roofline']['frac'], 4), 'e2e', round(d['e2e']['value'], 1), 'full', round(d['e2e_full_upload']['value'], 1), 'launches', d['gpu_launches'], 'c4', round(d['c4_strong']['value']), d['parity']['primary_ids']['ids_mismatch'])
P
