# A/B runs on the GPU box (scratch helper)
run() { # name lib
  NAUTILUS_CSM_LIB=$2 python bench.py --steps 5 --no-cpu-baseline > gpurun_out/ab_$1.json 2> gpurun_out/ab_$1.err
  python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/ab_$1.json").read().strip().splitlines()[-1])
    print("$1", round(l["ms_per_step"],3), {k:round(v["ms_per_step"],3) for k,v in l["kernels"].items()}, round(l["roofline"]["frac"],4))
except Exception as e:
    print("$1 failed", e, open("gpurun_out/ab_$1.err").read()[-400:])
PY
}
run main ""
run w14b2 nautilus_b200/lib/variants/w14b2.so
run w15b2 nautilus_b200/lib/variants/w15b2.so
run w16b2 nautilus_b200/lib/variants/w16b2.so
