#!/bin/bash
# Where does the epilogue warp spend its time?  clock64 counters (GABRIL_TC_DEBUG bit 2) with stores
# (bit 3) / bias loads (bit 4) / TMA (bit 0) switched off in turn.
out=gpurun_out/exp_epi.log
: > $out
for layer in conv1 conv4 conv5 res1x1; do
  for dbg in 4 12 20 28 5; do
    GABRIL_TC_DEBUG=$dbg GABRIL_TC_MSUB=1 python tools/one_layer.py $layer fwd 1024 5 >> $out 2>&1
  done
done
for dbg in 4 12; do
  GABRIL_TC_DEBUG=$dbg GABRIL_TC_MSUB=1 python tools/one_layer.py conv2 dgrad 1024 5 >> $out 2>&1
done
