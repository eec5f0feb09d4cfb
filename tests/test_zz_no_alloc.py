"""pfc_dot_noalloc (csrc/pfc_noalloc.cu): the reduction behind pfc_residual_norm with its scratch in a static __device__
array, for callers that cannot call cudaMalloc.  No counterpart in the reference.
Collected last on purpose (file name).

Checked through the C ABI, on two calls in a row that share the static scratch: sqrt(rhs . rhs) of the assembled
right-hand side against the existing entry point pfc_residual_norm (system_rhs.l2_norm(), mpi.cpp:584), a dot product of
two different vectors against numpy, no change of the free device memory across the calls, and the refusal of sizes
beyond PFC_NOALLOC_MAX_N."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

A = 4 * np.pi / np.sqrt(3.0)
P1, P2 = (-6 * A, -5 * A), (6 * A, 5 * A)
DT, EPS, PSI0 = 0.1, -0.8, -0.43
TOL = 1e-12  # relative; the two reductions sum the same numbers in different orders


def test_dot_noalloc_matches_residual_norm_twice(built):
    import torch
    P = built
    degree, n_refines = 1, 8                                   # 3 N = 198 147 doubles: 13 chunks, the last one ragged
    st = P.StandIn(degree, n_refines, P1, P2)
    Psi = st.hexagonal_lattice(P.default_A0(PSI0, EPS), PSI0, [[2 * A, 0], [-2 * A, 0]], [[A, 0], [-A, 0]])
    s = P.Solver(degree=degree, dt=DT, theta=1.0, eps=EPS, operator_kind=1, preconditioner=1)
    s.setup(st)
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    zero = torch.zeros(1, dtype=torch.float64, device="cuda")
    s.dot_noalloc(zero.data_ptr(), zero.data_ptr(), 0, out.data_ptr())   # warm-up: resolves the scratch address; n = 0 -> 0
    torch.cuda.synchronize()
    assert out.item() == 0.0
    rng = np.random.default_rng(4)
    states = [Psi, Psi + 1e-2 * rng.standard_normal(Psi.size)]
    for call, state in enumerate(states):                      # two calls in a row on the shared scratch
        s.set_state(state, Psi)
        s.assemble()
        ref = s.residual_norm()                                # existing entry point
        rhs = s.get_rhs()
        x = torch.from_numpy(rhs).cuda()
        y = torch.from_numpy(np.cos(np.arange(rhs.size) * 0.37)).cuda()
        out.fill_(float("nan"))
        out2 = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        s.dot_noalloc(x.data_ptr(), x.data_ptr(), rhs.size, out.data_ptr())
        s.dot_noalloc(x.data_ptr(), y.data_ptr(), rhs.size, out2.data_ptr())
        torch.cuda.synchronize()
        assert torch.cuda.mem_get_info()[0] >= free_before     # nothing was allocated on the device by the calls
        got = float(np.sqrt(out.item()))
        print(f"call {call}: pfc_residual_norm {ref:.16e}  sqrt(pfc_dot_noalloc) {got:.16e}")
        assert ref > 0 and abs(got - ref) <= TOL * ref
        want = float(np.dot(rhs, y.cpu().numpy()))
        assert abs(out2.item() - want) <= TOL * np.linalg.norm(rhs) * np.sqrt(rhs.size)
    # sizes beyond the static scratch are refused with a clear error (the pointers are never dereferenced)
    with pytest.raises(P.PfcError) as e:
        s.dot_noalloc(out.data_ptr(), out.data_ptr(), P.NOALLOC_MAX_N + 1, out.data_ptr())
    assert "PFC_NOALLOC_MAX_N" in str(e.value)
    s.close()
    st.close()
