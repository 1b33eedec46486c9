# grid-size sweep of k_generate / k_accumulate (blocks per SM) on book-one 4K
mkdir -p gpurun_out/$1
for cfg in "4 2" "4 8" "4 16" "8 8" "8 16"; do set -- $cfg $1
RT_B200_GEN_BLOCKS=$1 RT_B200_ACC_BLOCKS=$2 timeout 400 python bench.py --workload book-one-4k --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/$3/grid_$1_$2.json 2>gpurun_out/$3/grid_$1_$2.err
python - <<PY
import json
d=json.loads(open("gpurun_out/$3/grid_$1_$2.json").read().strip().splitlines()[-1])
print("book-one-4k gen_blocks $1 acc_blocks $2 value", round(d["value"],1), {k:round(v,1) for k,v in d["stage_ms_per_step"].items()})
PY
set -- $3
done
