#!/bin/bash
# tuning aid: time the K node-stencil variants (PFC_KS_VARIANT) with profiles/legs.py
for v in "$@"; do
  echo -n "variant $v: "
  PFC_KS_VARIANT=$v python profiles/legs.py --sweeps 4 --steps 1 2>&1 | tail -1 | python -c "import sys, json; d = json.loads(sys.stdin.read()); print(d['us_kmv'], d['us_precond_apply'], d['ms_per_step'])"
done
