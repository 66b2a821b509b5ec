for so in "$@"; do
  SRSRAN_LDPC_CUDA_LIB=$PWD/$so python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-pusch-extra 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$so', 'fixed6 %.2f' % d['value'], 'e2e %.2f Gb/s' % d['e2e']['value'])
"
done
