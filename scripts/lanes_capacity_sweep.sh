mkdir -p gpurun_out/r2f
for L in 2 4; do for cap in 262144 524288 1048576 2097152 4194304; do
RT_B200_LANES=$L timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --queue-capacity $cap > gpurun_out/r2f/c_L${L}_$cap.json 2>gpurun_out/r2f/c_L${L}_$cap.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2f/c_L${L}_$cap.json").read().strip().splitlines()[-1])
print("cornell lanes $L cap $cap value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1))
PY
done; done
for L in 3 4; do
RT_B200_LANES=$L timeout 300 python bench.py --workload book-one-4k --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2f/4k_L$L.json 2>gpurun_out/r2f/4k_L$L.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r2f/4k_L$L.json").read().strip().splitlines()[-1])
print("4k lanes $L value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1))
PY
done
