"""Multi-GPU parity worker (launched by torchrun from tests/test_gpu_multi.py, one rank per GPU).

Every rank builds its Morton chunk + ghost layer with the stand-in, runs `steps` timesteps through
the C ABI with NCCL halo exchange / all-reduce, and the owned entries of all ranks are summed into
the global state, which rank 0 compares with the CPU oracle (psi relative L2 <= 1e-9 per step).
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import pfc_b200 as P  # noqa: E402

A = P.LATTICE_A


def main():
    degree, n_refines, steps = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    kw = json.loads(sys.argv[4]) if len(sys.argv) > 4 else {}
    check_stress = kw.pop("check_stress", 0)
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    p1, p2 = (-6 * A, -5 * A), (6 * A, 5 * A)
    mesh = P.StandIn(degree, n_refines, p1, p2, rank=rank, n_ranks=world)
    s = P.Solver(degree=degree, device=local, **kw)
    uid = [P.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    s.comm_init(uid[0], rank, world)
    s.setup(mesh)
    if rank == 0:
        print(f"comm mode {s.comm_mode()} (1 NCCL, 2 peer memory)", flush=True)
    if os.environ.get("PFC_REQUIRE_P2P") == "1" and os.environ.get("PFC_NO_P2P", "0") == "0":
        assert s.comm_mode() == 2, "peer-memory path was not selected"
    Psi0 = mesh.hexagonal_lattice(P.default_A0(-0.43, -0.8), -0.43, [[2 * A, 0], [-2 * A, 0]], [[A, 0], [-A, 0]])
    s.set_state(Psi0, Psi0)
    if rank == 0:
        import oracle as O
        om = O.Mesh(degree, n_refines, p1, p2)
        Pn, Pn1 = Psi0.copy(), Psi0.copy()
    ok = True
    for step in range(steps):
        st = s.step()
        mine = torch.from_numpy(s.get_state()).cuda()   # owned entries, zeros elsewhere
        dist.all_reduce(mine)
        got = mine.cpu().numpy()
        if rank == 0:
            ref = om.newton_step(Pn, Pn1, 0.1, kw.get("theta", 1.0), -0.8)
            N = om.N
            err = np.linalg.norm(got[:N] - Pn[:N]) / np.linalg.norm(Pn[:N])
            err_all = np.linalg.norm(got - Pn) / np.linalg.norm(Pn)
            print(f"[{world} ranks] Q{degree} 2^{n_refines} step {step}: gmres gpu {st['gmres_iters']} oracle {ref['gmres_iters']} "
                  f"rel L2 psi {err:.2e} all {err_all:.2e}", flush=True)
            ok = ok and st["status"] == 0 and err <= 1e-9 and err_all <= 1e-9
    # owned-range transfers: the ranks' owned slices tile the global state; uploading them again (ghosts fetched from
    # their owners) reproduces the next step bit for bit
    own = s.get_state_owned()
    no = s.n_owned_nodes()
    counts = [None] * world
    dist.all_gather_object(counts, no)
    o0 = sum(counts[:rank])
    N_all = sum(counts)
    mine = s.get_state()
    same = all(np.array_equal(own[c * no:(c + 1) * no], mine[c * N_all + o0:c * N_all + o0 + no]) for c in range(3))
    s.set_state_owned(own, own)
    full = torch.from_numpy(s.get_state()).cuda()
    dist.all_reduce(full)
    st_a = s.step()
    after_a = s.get_state_owned().copy()
    s.set_state(full.cpu().numpy(), full.cpu().numpy())
    st_b = s.step()
    same = same and np.array_equal(after_a, s.get_state_owned()) and st_a["gmres_iters"] == st_b["gmres_iters"]
    flag_o = torch.tensor([1 if same else 0], device="cuda")
    dist.all_reduce(flag_o, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"[{world} ranks] owned-range transfers consistent: {bool(flag_o.item())}", flush=True)
        ok = ok and bool(flag_o.item())
        ref = om.newton_step(Pn, Pn1, 0.1, kw.get("theta", 1.0), -0.8)  # keep the oracle in step with the extra step
    if check_stress:  # stress projection of the stepped state, owned nodes summed over the ranks
        sg, its = s.stress(-0.8)
        so, its_o = s.stress(-0.8, owned=True)  # the rank's slice of the same result
        assert its_o == its and all(np.array_equal(so[c], sg[c, o0:o0 + no]) for c in range(4))
        mine = torch.from_numpy(sg).cuda()
        dist.all_reduce(mine)
        sg = mine.cpu().numpy()
        if rank == 0:
            o = om.stress(Pn, -0.8)
            err = np.abs(sg - o["stress"]).max() / np.abs(o["stress"]).max()
            print(f"[{world} ranks] stress: CG its gpu {its} oracle {o['cg_iters']} max rel diff {err:.2e}", flush=True)
            ok = ok and o["status"] == 0 and err <= 1e-8 and all(abs(a - b) <= 1 for a, b in zip(its, o["cg_iters"]))
        # every rank writes its owned cells (ghost vertex values included), rank 0 the record
        import tempfile
        out = [tempfile.mkdtemp() + os.sep if rank == 0 else None]
        dist.broadcast_object_list(out, src=0)
        mine = torch.from_numpy(s.get_state()).cuda()
        dist.all_reduce(mine)
        got = mine.cpu().numpy()
        s.output(mesh, 0, out[0], "hexagonal_lattice", steps)
        s.output(mesh, 2, out[0], "stress", 0)
        dist.barrier()
        if rank == 0:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            from vtu_reader import read_piece, read_record
            key = {tuple(np.round(p, 9)): t for t, p in enumerate(om.support)}
            rec = read_record(out[0] + f"hexagonal_lattice_{steps:02d}.pvtu")
            ok = ok and len(rec["pieces"]) == world
            n_cells, worst = 0, 0.0
            for q in range(world):
                pc = read_piece(out[0] + rec["pieces"][q])
                ps = read_piece(out[0] + f"stress_00.{q}.vtu")
                node = np.array([key[tuple(np.round(p[:2], 9))] for p in pc["points"]])
                n_cells += pc["n_cells"]
                for c, a in enumerate(("psi", "chi", "phi")):
                    worst = max(worst, np.abs(pc["point_data"][a] - got[c * om.N + node]).max())
                for c, slot in enumerate((0, 1, 3, 4)):
                    worst = max(worst, np.abs(ps["point_data"]["stress"][:, slot] - sg[c, node]).max())
            print(f"[{world} ranks] output: {n_cells} cells in {world} pieces, max |written - state| {worst:.2e}", flush=True)
            ok = ok and n_cells == 4 ** n_refines and worst == 0.0
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, src=0)
    s.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
