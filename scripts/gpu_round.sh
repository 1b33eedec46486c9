#!/bin/bash
# One GPU-box visit: tests, smoke, bench (both arms), per-scene stage timing, ncu launch list + full captures.
#   gpurun --timeout 1500 -- 'bash scripts/gpu_round.sh <tag> [steps...]'     steps default: test smoke bench ref scenes launches full
cd "$(dirname "$0")/.."
TAG=${1:-r1}; shift
STEPS=${*:-test smoke bench ref scenes launches full}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/gpu.txt 2>&1
for s in $STEPS; do
  case $s in
    test)   timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log; tail -5 $OUT/pytest_gpu.log ;;
    smoke)  timeout 300 python __graft_entry__.py smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke.log ;;
    bench)  timeout 600 python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; cat $OUT/bench.json ;;
    ref)    timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref rc=$?"; cat $OUT/bench_ref.json ;;
    scenes) timeout 600 python scripts/time_scenes.py > $OUT/scenes.log 2>&1; echo "scenes rc=$?"; cat $OUT/scenes.log ;;
    variants)
      RT_B200_LIB=$PWD/ray-tracer_b200/librt_b200.so python scripts/time_scenes.py $VSCENES > $OUT/variants.log 2>&1
      bash scripts/bench_variants.sh $VSCENES >> $OUT/variants.log 2>&1; echo "variants rc=$?"; cat $OUT/variants.log ;;
    allbench)
      for wl in book-one book-two stress-1m book-one-4k; do
        timeout 900 python bench.py --workload $wl --steps 2 --warmup 3 --no-cpu-baseline > $OUT/bench_$wl.json 2> $OUT/bench_$wl.err; echo "bench $wl rc=$?"; cat $OUT/bench_$wl.json
      done ;;
    launches)
      timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-book-one-4k > $OUT/plain_launches.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 4000 --csv --log-file $OUT/launches.csv \
        python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $OUT/ncu_launches.log 2>&1
      echo "launches rc=$?" ;;
    full)
      IFS=';' read -ra SCS <<< "${FULLSCENES:-cornell-box 600 600 16 50;book-one 1920 1080 4 50}"
      for sc in "${SCS[@]}"; do
        set -- $sc
        timeout 120 python scripts/profile_target.py $sc > $OUT/plain_$1.log 2>&1 &&
        timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${NCU_REGEX:-k_extend|k_shade}" -s ${NCU_SKIP:-40} -c ${NCU_COUNT:-4} -o $OUT/full_$1 -f \
          python scripts/profile_target.py $sc > $OUT/ncu_full_$1.log 2>&1
        echo "full $1 rc=$?"
      done ;;
  esac
done
ls -la $OUT
