#!/bin/bash
# tuning aid: rows per strip of the streaming kernels (PFC_RING_RS; unset = the wave-aware choice of ring_rows_per_strip)
for nr in ${NRS:-12 11 10}; do
  for rs in ${RSS:-auto 16 32 48 64 80 128 240 320}; do
    echo "== Q1 n_refines $nr rs=$rs"
    if [ "$rs" = "auto" ]; then unset PFC_RING_RS; else export PFC_RING_RS=$rs; fi
    python profiles/legs.py --degree 1 --n-refines $nr --sweeps 4 --steps 1 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print({k: (round(v, 1) if isinstance(v, float) else ([round(x, 1) for x in v] if isinstance(v, list) and v and isinstance(v[0], float) else v)) for k, v in d.items() if k.startswith('us_precond') or k in ('ms_per_step',)})
"
  done
done
