# A/B helper on the GPU box (scratch)
export NAUTILUS_CSM_LIB=$PWD/nautilus_b200/lib/variants/tune.so
for n in 32 64 128 250 500 1000; do
for tr in 0 1; do
NAUTILUS_CSM_TWO_RUN=$tr python bench.py --scans $n --steps 10 --no-cpu-baseline > gpurun_out/ab.json 2> gpurun_out/ab.err
python - <<PY
import json
l=json.loads(open("gpurun_out/ab.json").read().strip().splitlines()[-1])
print("scans=$n two_run=$tr", round(l["ms_per_step"],3), {k:round(v["ms_per_step"],3) for k,v in l["kernels"].items()})
PY
done
done
