#!/bin/bash
# time tuning builds of the whole-model kernel at cfg 2a (profiles/build_variants.py): bench.py main line only
for lib in "$@"; do
  CTRHOT_LIB=$lib python bench.py --steps 50 --warmup 5 --no-secondary --no-sharded --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.readline()); print('%-70s %8.2f us/step  kernel %.2f us  value %.1f M/s' % ('$lib'.split('/')[-1], d['ms_per_step']*1e3, d['roofline']['ms']*1e3, d['value']/1e6))"
done
