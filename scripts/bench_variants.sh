#!/bin/bash
# Run scripts/time_scenes.py against every library in ray-tracer_b200/build/variants (on the GPU box).
cd "$(dirname "$0")/.."
for so in ray-tracer_b200/build/variants/*.so; do
  RT_B200_LIB=$PWD/$so python scripts/time_scenes.py "$@"
done
