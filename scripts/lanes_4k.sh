#!/bin/bash
# book-one 4K (the north-star scene) under different numbers of wavefront lanes / queue capacities
for cfg in "2 0" "3 0" "4 0" "2 8388608" "3 2097152"; do
  set -- $cfg
  RT_B200_LANES=$1 python bench.py --workload book-one-4k --steps 3 --no-cpu-baseline --queue-capacity $2 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('lanes $1 cap $2:', round(d['value'],1), 'Mpaths/s', round(d['ms_per_step'],1), 'ms  e2e', round(d['e2e']['value'],1))"
done
