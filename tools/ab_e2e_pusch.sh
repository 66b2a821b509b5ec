#!/bin/bash
# Development helper: both host-fed figures (cfg 4 e2e, full-slot pusch e2e) once per library variant in build/ab/, same box.
for so in "$@"; do
  SRSRAN_LDPC_CUDA_LIB=$PWD/$so python bench.py --steps 4 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); p=d['pusch_slot']; print('$so', 'fixed6 %.2f' % d['value'], 'e2e %.2f' % d['e2e']['value'], 'pusch %.2f' % p['value'], 'pusch e2e %.2f Gb/s' % p['e2e']['value'])
    elif 'rror' in l: print(l.strip())
"
done
