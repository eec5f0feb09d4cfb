#!/bin/bash
# tuning aid: legs of the ring path for prefetch depth (PFC_RING_PF: 1 shallow, 2 deep) / rows per strip (PFC_RING_RS)
for wl in "--degree 1 --n-refines 12 --sweeps 3" "--degree 2 --n-refines 9 --sweeps 5"; do
  for pf in ${PFS:-1 2}; do for rs in ${RSS:-16 32 64 128}; do
    echo "== $wl pf=$pf rs=$rs"
    PFC_RING_PF=$pf PFC_RING_RS=$rs python profiles/legs.py $wl --steps 1 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print({k: (round(v, 1) if isinstance(v, float) else ([round(x, 1) for x in v] if isinstance(v, list) and v and isinstance(v[0], float) else v)) for k, v in d.items() if k.startswith('us_precond') or k in ('ms_per_step', 'gmres_iters')})
"
  done; done
done
