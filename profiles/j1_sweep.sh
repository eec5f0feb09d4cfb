#!/bin/bash
# tuning aid: variants of the Q1 J apply kernel (PFC_J1_VARIANT, csrc/pfc_lattice.cu): self-check, then operator time on M4
for v in ${VARIANTS:-0 1 2 3 4 5 6 7 8}; do
  echo "== variant $v"
  PFC_J1_VARIANT=$v python -m pytest tests/test_gpu_parity.py -x -q -k "ring_smoother" 2>&1 | tail -1
  PFC_J1_VARIANT=$v python profiles/legs.py --degree 1 --n-refines 12 --sweeps 3 --steps 1 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print({k: round(v, 1) for k, v in d.items() if k in ('us_operator', 'gbs_operator', 'ms_per_step', 'us_precond_apply')})
"
done
