"""N > 1 host logic with real message passing on CPU: world_size 2 (and 4) over gloo, see tests/gloo_worker.py."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("degree,n_refines,world,flags", [(1, 4, 2, 0), (2, 3, 2, 0), (1, 4, 2, 1), (2, 3, 2, 3), (1, 4, 4, 1)])
def test_ranks_over_gloo(built, degree, n_refines, world, flags):
    """flags: 0 = SELL numbering, 1 = lattice numbering with tables, 3 = lattice numbering without tables (halo only)."""
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29541", os.path.join(ROOT, "tests", "gloo_worker.py"),
           str(degree), str(n_refines), str(flags)]
    env = dict(os.environ, OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    print(r.stdout[-2000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
