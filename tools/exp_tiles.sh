#!/bin/bash
# Tile-shape experiments for the tcgen05 kernel (one gpurun call): accumulator ring depth, 256-row
# CTA tiles, and which side (TMA feed / MMA+epilogue) bounds each layer.
B=${1:-1024}
out=gpurun_out/exp_tiles_b$B.log
: > $out
run() { echo "=== $*" >> $out; env "$@" python tools/bench_layers.py $B >> $out 2>&1; }
run GABRIL_TC_MSUB=1 GABRIL_TC_ACC=2
run GABRIL_TC_MSUB=1 GABRIL_TC_ACC=8
run GABRIL_TC_MSUB=2 GABRIL_TC_ACC=8
run GABRIL_TC_MSUB=2 GABRIL_TC_ACC=2
run GABRIL_TC_MSUB=1 GABRIL_TC_ACC=2 GABRIL_TC_OCC=1
run GABRIL_TC_MSUB=1 GABRIL_TC_ACC=2 GABRIL_TC_DEBUG=1
run GABRIL_TC_MSUB=1 GABRIL_TC_ACC=2 GABRIL_TC_DEBUG=2
run GABRIL_TC_MSUB=2 GABRIL_TC_ACC=8 GABRIL_TC_DEBUG=1
run GABRIL_TC_MSUB=2 GABRIL_TC_ACC=8 GABRIL_TC_DEBUG=2
