#!/bin/bash
# config 5 at G = 1, fused exchange only, for each cluster size of the whole-model kernel
for c in 1 2 4; do
  echo "cluster=$c"; CTRHOT_CFG5_CLUSTER=$c timeout 200 python profiles/bench_cfg5.py 10000000 fused
done
