#!/bin/bash
# development aid: time the symmetric force kernel variants (NBODY_SYM_VARIANT) at N=262144
for v in ${VARIANTS:-0 11 13 14 15}; do
    echo "sym variant $v: $(NBODY_SYM_VARIANT=$v python tools/quick_bench.py ${CASE:-262144:1:5} 2>&1 | grep 'n=')"
done
