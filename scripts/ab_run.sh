run() { # name lib stagecells
  NAUTILUS_CSM_LIB=$2 python bench.py --steps 5 --no-cpu-baseline --stage-cells $3 > gpurun_out/ab_$1.json 2> gpurun_out/ab_$1.err
  python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/ab_$1.json").read().strip().splitlines()[-1])
    print("$1", round(l["ms_per_step"],3), {k:round(v["ms_per_step"],3) for k,v in l["kernels"].items()}, round(l["roofline"]["frac"],4))
except Exception as e:
    print("$1 failed", e, open("gpurun_out/ab_$1.err").read()[-400:])
PY
}
python -m pytest tests/test_gpu_parity.py tests/test_gpu_wide.py -x -q -m gpu 2>&1 | tail -2
run main "" 0
run w10b3 nautilus_b200/lib/variants/w10b3.so 8600
run w10b3_d nautilus_b200/lib/variants/w10b3.so 0
run w16b2 nautilus_b200/lib/variants/w16b2.so 0
run w8b4 nautilus_b200/lib/variants/w8b4.so 6400
