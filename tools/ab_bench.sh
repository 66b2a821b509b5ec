#!/bin/bash
# Development helper: runs bench.py (kernel-only, no CPU baseline / e2e) once per library variant in build/ab/, same box.
for so in "$@"; do
  SRSRAN_LDPC_CUDA_LIB=$PWD/$so python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-pusch-extra 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$so', 'fixed6 %.2f Gb/s' % d['value'], 'early-stop %.2f Gb/s' % d['early_stop']['value'], d['clocks'])
    elif 'rror' in l: print(l.strip())
"
done
