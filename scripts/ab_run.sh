# A/B helper on the GPU box (scratch)
python -m pytest tests/test_gpu_parity.py tests/test_gpu_solver_config.py tests/test_gpu_soak.py tests/test_gpu_wide.py -x -q -m gpu 2>&1 | tail -3
for n in 2000 400 250; do
python bench.py --scans $n --steps 10 --no-cpu-baseline > gpurun_out/ab.json 2> gpurun_out/ab.err
python - <<PY
import json
l=json.loads(open("gpurun_out/ab.json").read().strip().splitlines()[-1])
print("scans=$n", round(l["ms_per_step"],3), {k:round(v["ms_per_step"],3) for k,v in l["kernels"].items()})
PY
done
python scripts/bench_configs.py solver wide > gpurun_out/ab_solver.json 2> gpurun_out/ab_solver.err
python - <<PY
import json
for line in open("gpurun_out/ab_solver.json"):
    if line.startswith("{"):
        l=json.loads(line); print(l["config"][:30], round(l["ms"],3), l["kernels_ms"], l["pairs_per_sec"])
PY
