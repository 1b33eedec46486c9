bash scripts/gpu_round.sh r2o test
for v in 0 1; do
RT_B200_SHADE_RAYS2=$v timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r2o/cornell_rays2_$v.json 2>gpurun_out/r2o/cornell_rays2_$v.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2o/cornell_rays2_$v.json").read().strip().splitlines()[-1])
print("cornell shade_rays2=$v value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), {k:round(v,1) for k,v in d["stage_ms_per_step"].items()})
PY
done
