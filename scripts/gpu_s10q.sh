mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x --timeout=120 > gpurun_out/s10q_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/s10q_pytest.log
( time timeout 900 python bench.py > gpurun_out/s10q_bench.json 2> gpurun_out/s10q_bench.err ) 2>&1 | grep real; python - <<'PY'
import json
d=json.load(open('gpurun_out/s10q_bench.json'))
print({k: d[k] for k in ("value","ms_per_step","gpu_launches")}, d["e2e"]["value"], d["mpc_step"], d["pinv_build"]["ms"])
PY
tail -3 gpurun_out/s10q_bench.err
