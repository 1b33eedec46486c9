# bench.py over a set of workloads:  bash scripts/bench_set.sh <tag> <steps> wl1 wl2 ...
TAG=$1; STEPS=$2; shift 2
mkdir -p gpurun_out/$TAG
for wl in "$@"; do
timeout 600 python bench.py --workload $wl --steps $STEPS --warmup 3 --no-cpu-baseline > gpurun_out/$TAG/bench_$wl.json 2>gpurun_out/$TAG/bench_$wl.err
python - <<PY
import json
d=json.loads(open("gpurun_out/$TAG/bench_$wl.json").read().strip().splitlines()[-1])
print("$wl value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), {k:round(v,1) for k,v in d["stage_ms_per_step"].items()}, "fp32 frac", round(d["roofline_fp32"]["frac"],3))
PY
done
