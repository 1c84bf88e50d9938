# A/B helper on the GPU box (scratch)
python -m pytest tests/test_gpu_log.py tests/test_gpu_parity.py tests/test_gpu_solver_config.py tests/test_gpu_soak.py tests/test_gpu_wide.py tests/test_gpu_uncertainty.py -x -q -m gpu 2>&1 | tail -5
python bench.py --steps 5 --no-cpu-baseline > gpurun_out/ab_0.json 2> gpurun_out/ab_0.err
python - <<PY
import json
l=json.loads(open("gpurun_out/ab_0.json").read().strip().splitlines()[-1])
print("bench", round(l["ms_per_step"],3), {k:round(v["ms_per_step"],3) for k,v in l["kernels"].items()})
PY
python scripts/bench_configs.py solver > gpurun_out/ab_solver.json 2> gpurun_out/ab_solver.err
python - <<PY
import json
l=json.loads(open("gpurun_out/ab_solver.json").read().strip().splitlines()[-1])
print("solver", round(l["ms"],3), l["kernels_ms"], l["pairs_per_sec"])
PY
