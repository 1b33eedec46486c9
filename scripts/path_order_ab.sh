# A/B of the path issue order (RT_B200_PATH_ORDER 0 = sample-major, 1 = pixel-major) on every bench workload
mkdir -p gpurun_out/$1
for wl in cornell-box book-one-4k stress-1m book-two; do for po in 0 1; do
RT_B200_PATH_ORDER=$po timeout 400 python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/$1/po${po}_$wl.json 2>gpurun_out/$1/po${po}_$wl.err
python - <<PY
import json
d=json.loads(open("gpurun_out/$1/po${po}_$wl.json").read().strip().splitlines()[-1])
print("$wl order $po value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), {k:round(v,1) for k,v in d["stage_ms_per_step"].items()})
PY
done; done
