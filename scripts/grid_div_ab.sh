# experiment: half-size extend / shade grids so that the kernels of two lanes co-reside on every SM
mkdir -p gpurun_out/$1
for wl in cornell-box book-one-4k; do for cfg in "1 2" "2 2" "2 3" "2 4"; do set -- $cfg $1 $wl
RT_B200_GRID_DIV=$1 RT_B200_LANES=$2 timeout 400 python bench.py --workload $4 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/$3/gd$1_L$2_$4.json 2>gpurun_out/$3/gd$1_L$2_$4.err
python - <<PY
import json
d=json.loads(open("gpurun_out/$3/gd$1_L$2_$4.json").read().strip().splitlines()[-1])
print("$4 grid_div $1 lanes $2 value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1))
PY
set -- $3
done; done
