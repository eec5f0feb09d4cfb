#!/bin/bash
# tuning aid: Gram-Schmidt bandwidth against the distance between the Krylov basis vectors (PFC_STRIDE_PAD, doubles)
for pad in ${PADS:-0 16 64 256 1024 4096 16384 65536 262144 1048576 1234576 3000016}; do
  echo "== pad $pad"
  PFC_STRIDE_PAD=$pad DIMS="${DIMS:-12 16 20 24 27}" python profiles/krylov_dims.py 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print(d['dim'], round(d['gbs_dot']), round(d['gbs_axpy']))
"
done
