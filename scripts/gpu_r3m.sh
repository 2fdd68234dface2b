#!/bin/bash
# r3m: last confirmation of the committed tree: GPU suite, smoke(), default bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/r3m_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r3m_pytest.txt; tail -n 3 gpurun_out/r3m_pytest.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r3m_smoke.txt 2>&1; tail -n 2 gpurun_out/r3m_smoke.txt
timeout 900 python bench.py > gpurun_out/r3m_bench.json 2> gpurun_out/r3m_bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r3m_bench.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"], d["roofline"]["frac"], d["clocks"], {k:round(v["value"],1) for k,v in d["other_configs"].items()})
PY
