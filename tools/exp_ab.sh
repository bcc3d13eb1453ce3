#!/bin/bash
# A/B of two builds of the library on the same box: flat-slab layers, im2col layers, the traced step.
OLD=$PWD/gabril-carla_b200/libgabril_b200_old.so
out=gpurun_out/exp_ab.log
: > $out
for lib in "" "$OLD"; do
  echo "=== lib=${lib:-new}" >> $out
  GABRIL_B200_LIB=$lib GABRIL_TC_MSUB=1 python tools/bench_flat.py 1024 >> $out 2>&1
  GABRIL_B200_LIB=$lib GABRIL_TC_MSUB=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-store-feed --trace-out gpurun_out/ab_trace_$( [ -z "$lib" ] && echo new || echo old ).csv 2>&1 | cut -c1-250 >> $out
done
for l in conv4 conv1; do GABRIL_TC_DEBUG=4 GABRIL_TC_MSUB=1 python tools/one_layer.py $l fwd 1024 5 >> $out 2>&1; done
