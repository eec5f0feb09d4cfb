"""Multi-GPU parity: N ranks over NCCL == oracle.  Needs >= 2 GPUs (skipped otherwise)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("degree,n_refines,steps,extra", [(1, 5, 2, "{}"), (2, 4, 1, "{}"), (1, 4, 1, '{"use_cuda_graph": 0}'),
                                                          (1, 5, 2, '{"preconditioner": 1}'), (2, 4, 1, '{"preconditioner": 1}'),
                                                          (1, 5, 2, '{"operator_kind": 1}'), (2, 4, 2, '{"operator_kind": 1, "preconditioner": 1}'),
                                                          (1, 5, 2, '{"operator_kind": 1, "preconditioner": 1}'),
                                                          (2, 4, 1, '{"operator_kind": 1, "preconditioner": 1, "check_stress": 1}'),
                                                          (1, 6, 2, '{"operator_kind": 1, "preconditioner": 2}'),
                                                          (1, 5, 2, '{"operator_kind": 1, "preconditioner": 2, "theta": 0.5}')])
@pytest.mark.parametrize("no_p2p", [0, 1])
def test_ranks_match_oracle(built, degree, n_refines, steps, extra, no_p2p):
    """no_p2p = 0: halo exchange and all-reduce by the peer-memory kernels over NVLink; 1: NCCL."""
    if no_p2p and "operator_kind" not in extra and "preconditioner" not in extra and steps > 1:
        pytest.skip("NCCL path covered by the other cases")
    n = _n_gpus()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    # 2 ranks for every case; the full box (up to 8 ranks) only for the bench-default configuration over peer memory,
    # so that the suite stays within minutes on an 8-GPU box
    worlds = {2}
    if ('"operator_kind": 1, "preconditioner": 1' in extra or '"preconditioner": 2}' in extra) and not no_p2p:
        worlds.add(min(n, 8))
    for world in sorted(worlds):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", "29517", os.path.join(ROOT, "tests", "mgpu_worker.py"),
               str(degree), str(n_refines), str(steps), extra]
        env = dict(os.environ, PFC_NO_P2P=str(no_p2p))
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
        print(r.stdout[-3000:])
        assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
