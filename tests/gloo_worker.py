"""world_size-2 CPU worker (gloo) for the host side of the N > 1 path (launched by tests/test_multi_gloo_cpu.py).

Each process plays one rank: stand-in chunk + host plan (what pfc_setup derives), then with REAL message passing
  1. the send-list negotiation of comm_exchange_send_lists (csrc/pfc_comm.cu): counts all-gathered, wanted global
     nodes sent to their owners, owners translate them to internal rows through iperm;
  2. a halo exchange of a vector in internal order (pack by send_rows, receive into the ghost tail);
  3. the rank-local operator (numpy emulation of the SELL kernel, tests/emul.py) and an all-reduced dot product.
Rank 0 compares y = J x, -R and the dot product with the single-rank oracle.  No GPU, no product arithmetic here.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import emul  # noqa: E402
import pfc_b200 as P  # noqa: E402

A = P.LATTICE_A
DT, THETA, EPS = 0.1, 0.5, -0.8


def main():
    degree, n_refines, flags = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    p1, p2 = (-6 * A, -5 * A), (6 * A, 5 * A)
    st = P.StandIn(degree, n_refines, p1, p2, rank=rank, n_ranks=world)
    plan = P.HostPlan(st.mesh_ptr, degree, flags)
    N = st.N
    if flags & 1:
        assert plan.lattice["active"] == 1, plan.lattice["why"]

    # ---- 1. send lists ----
    counts = torch.zeros(world, dtype=torch.int64)
    for k, peer in enumerate(plan.peers):
        counts[peer] = int(plan.recv_ptr[k + 1] - plan.recv_ptr[k])
    allc = [torch.zeros(world, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(allc, counts)
    allc = torch.stack(allc).numpy()                     # allc[r][o] = nodes rank r needs from owner o
    peers = [r for r in range(world) if r != rank and (allc[rank][r] > 0 or allc[r][rank] > 0)]
    assert list(plan.peers) == [r for r in peers if allc[rank][r] > 0]
    reqs, asked = [], {}
    for k, peer in enumerate(plan.peers):
        want = torch.from_numpy(np.ascontiguousarray(plan.ghost_global[plan.recv_ptr[k]:plan.recv_ptr[k + 1]]).astype(np.int64))
        reqs.append(dist.isend(want, dst=int(peer)))
    for r in peers:
        if allc[r][rank] > 0:
            buf = torch.zeros(int(allc[r][rank]), dtype=torch.int64)
            dist.recv(buf, src=r)
            asked[r] = buf.numpy()
    for q in reqs:
        q.wait()
    send_rows = {}
    for r, nodes in asked.items():
        assert np.all((nodes >= plan.own0) & (nodes < plan.own0 + plan.n_owned)), "asked for a node this rank does not own"
        send_rows[r] = plan.iperm[nodes - plan.own0]

    # ---- 2. halo exchange of an internal vector ----
    rng = np.random.default_rng(11)
    x_api = rng.standard_normal(3 * N)                   # same on every rank
    sl, ms = st.constraint_slave, st.constraint_master
    x_api[sl] = x_api[ms]
    x = emul.to_internal(plan, x_api, N)
    ref_ghost = x[3 * plan.Np:].copy()
    x[3 * plan.Np:] = np.nan                              # must come from the owners
    reqs = []
    for r, rows in send_rows.items():
        buf = np.stack([x[c * plan.Np + rows] for c in range(3)], axis=1).reshape(-1)   # [k][c] as halo_pack_kernel
        reqs.append(dist.isend(torch.from_numpy(buf.copy()), dst=r))
    for k, peer in enumerate(plan.peers):
        n = int(plan.recv_ptr[k + 1] - plan.recv_ptr[k])
        buf = torch.zeros(3 * n, dtype=torch.float64)
        dist.recv(buf, src=int(peer))
        x[3 * plan.Np + 3 * int(plan.recv_ptr[k]):3 * plan.Np + 3 * int(plan.recv_ptr[k + 1])] = buf.numpy()
    for q in reqs:
        q.wait()
    assert np.array_equal(x[3 * plan.Np:], ref_ghost), "halo exchange does not reproduce the owners' values"

    ok = True
    if not (flags & 2):
        # ---- 3. rank-local assembly + operator, all-reduced results ----
        Psi = st.hexagonal_lattice(P.default_A0(-0.43, EPS), -0.43, [[2 * A, 0], [-2 * A, 0]], [[A, 0], [-A, 0]])
        Psi1 = Psi + 1e-2 * np.random.default_rng(7).standard_normal(Psi.size)
        Psi1[sl] = Psi1[ms]
        tab = (st.shape, st.dshape, st.qw)
        pn, pn1 = emul.to_internal(plan, Psi1, N), emul.to_internal(plan, Psi, N)
        B, C, M, K, rhs = emul.local_blocks(plan, tab, pn, pn1, DT, THETA, EPS)
        vB, vC, vM, vK, rr = emul.scatter(plan, B, C, M, K, rhs)
        y = emul.spmv_fast(plan, vB, vC, vM, vK, x, DT * THETA)
        got_y = torch.from_numpy(emul.owned_to_api(plan, y, N, np.zeros(3 * N)))
        got_r = torch.from_numpy(emul.owned_to_api(plan, rr, N, np.zeros(3 * N)))
        dot = torch.tensor([float(np.dot(y[:3 * plan.Np], x[:3 * plan.Np]))], dtype=torch.float64)
        dist.all_reduce(got_y)
        dist.all_reduce(got_r)
        dist.all_reduce(dot)
        if rank == 0:
            import oracle as O
            om = O.Mesh(degree, n_refines, p1, p2)
            ovals, orhs = om.assemble(Psi1, Psi, DT, THETA, EPS)
            ref = om.spmv(ovals, x_api)
            e1 = np.max(np.abs(got_y.numpy() - ref)) / np.max(np.abs(ref))
            e2 = np.max(np.abs(got_r.numpy() - orhs)) / np.max(np.abs(orhs))
            e3 = abs(dot.item() - float(np.dot(ref, x_api))) / abs(float(np.dot(ref, x_api)))
            print(f"gloo {world} ranks Q{degree} 2^{n_refines}: y {e1:.2e} rhs {e2:.2e} dot {e3:.2e}", flush=True)
            ok = e1 <= 1e-12 and e2 <= 1e-12 and e3 <= 1e-12
    flag = torch.tensor([1 if ok else 0])
    dist.broadcast(flag, src=0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
