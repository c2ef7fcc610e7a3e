#!/usr/bin/env bash
# One GPU-box pass: parity tests, smoke, bench (both arms), launch list.  Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1
echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
timeout 120 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 400 python bench.py "$@" > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -3 gpurun_out/bench_n1.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_n1.json"))
print("value %.4g samples/s  %.3f us/step  frac %.3f  e2e %.4g  e2e_resident %.4g  launches %d  cpu1 %.4g" % (
    d["value"], 1e3 * d["ms_per_step"], d["roofline"]["frac"], d["e2e"]["value"],
    d.get("e2e_resident", {}).get("value", 0), d["gpu_launches"], d["cpu_baseline"]["value"]))
PY
timeout 300 python bench.py --impl reference --steps 200 --warmup 5 > gpurun_out/bench_ref.json 2>&1
python -c "import json;d=json.load(open('gpurun_out/bench_ref.json'));print('reference arm %.4g samples/s on %d cores' % (d['value'], d['cpu_baseline']['cores']))"
