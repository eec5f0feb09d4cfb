#!/bin/bash
# tuning aid: time the J / Schur-row cell kernels for the tile shapes of PFC_LAT_TILE with profiles/legs.py
for v in "$@"; do
  echo -n "tile $v: "
  PFC_LAT_TILE=$v python profiles/legs.py --sweeps 4 --steps 1 ${DEG:+--degree $DEG} ${NREF:+--n-refines $NREF} 2>&1 | tail -1 | python -c "import sys, json; d = json.loads(sys.stdin.read()); print('J', round(d['us_operator'],1), 'S', round(d['us_schur_row'],1), 'K', round(d['us_kmv'],1), 'precond', round(d['us_precond_apply'],1), 'step ms', round(d['ms_per_step'],2))"
done
