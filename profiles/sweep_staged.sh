#!/bin/bash
# tuning aid: Chebyshev sweep with operands staged by cp.async (PFC_SWEEP_STAGED = prefetch distance, 0 = register prefetch)
for v in ${VARIANTS:-0 3 4 6}; do
  echo "== staged $v"
  PFC_SWEEP_STAGED=$v python -m pytest tests/test_gpu_parity.py -x -q -k "ring_smoother or bench_size_step" 2>&1 | tail -1
  for nr in 12 11; do
  PFC_SWEEP_STAGED=$v python profiles/legs.py --degree 1 --n-refines $nr --sweeps 4 --steps 1 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print({k: (round(v, 1) if isinstance(v, float) else ([round(x, 1) for x in v] if isinstance(v, list) and v and isinstance(v[0], float) else v)) for k, v in d.items() if k.startswith('us_precond') or k in ('ms_per_step', 'gmres_iters')})
"
  done
done
