#!/bin/bash
# Development helper: demodulator line of bench.py once per library variant in build/ab/, same box (small decode workload).
for so in "$@"; do
  SRSRAN_LDPC_CUDA_LIB=$PWD/$so python bench.py --codeblocks 2048 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); m=d['demodulator']; print('$so', 'demod %.4f ms  %.0f GB/s' % (m['ms_per_step'], m['roofline']['achieved']))
    elif 'rror' in l: print(l.strip())
"
done
