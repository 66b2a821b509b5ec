#!/bin/bash
# Development helper: like ab_bench.sh but keeps the pusch_slot extra (fused dematch + decode) in the run.
for so in "$@"; do
  SRSRAN_LDPC_CUDA_LIB=$PWD/$so python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$so', 'fixed6 %.2f' % d['value'], 'early-stop %.2f' % d['early_stop']['value'], 'pusch_slot %.2f Gb/s' % d['pusch_slot']['value'], d['pusch_slot']['crc_pass_rate'], d['clocks']['sm_mhz'])
    elif 'rror' in l: print(l.strip())
"
done
