"""Run under torchrun with >= 2 GPUs:  python -m torch.distributed.run --nproc-per-node 2 tests/mg_nccl_check.py
Partitioned residual + tangent assembly with the NCCL exchange inside the library; every rank's owned rows are
compared bit for bit with the golden vectors of the unmodified reference. Prints one OK line per rank."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mimi_b200 import capi, partition as P  # noqa: E402
from tests.cases import MATS, load_golden  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok_all = True
    for name in ("cube_p2_n3_patch", "cube_p2_n2"):
        uid = [capi.mg_unique_id() if rank == 0 else None]   # one id per communicator
        dist.broadcast_object_list(uid, src=0)
        D = load_golden(name)
        dim = D["dim"]
        R = P.rank_problem(D["v_dofs"], dim, D["NV"], P.element_ranges(D["NE"], world), rank)
        ctx = capi.Context(local)
        loc = R["local_elems"]
        ctx.set_connectivity(dim, D["nd"], D["nq"], D["NV"], R["v_dofs"])
        ctx.upload_quad_tables(D["dN_dX"][loc], D["w"][loc], D["detJ"][loc])
        ctx.mg_setup(R)
        ctx.mg_init_nccl(rank, world, uid[0])
        ctx.build_pattern()
        ctx.set_material(MATS["neohookean"], list(D["lambda_mu"]))
        rows = np.repeat(R["owned"], dim)
        gI = D["I"].astype(np.int64)
        idx = np.concatenate([np.arange(gI[v], gI[v + 1]) for v in np.nonzero(rows)[0]])
        for rep in range(2):
            ctx.set_u(D["u"]); ctx.set_r(None); ctx.set_A(np.ascontiguousarray(D["mass"][idx]))
            ctx.assemble_residual_and_grad_device(D["gf"])
            ok = np.array_equal(ctx.get_r()[rows], D["r2_neohookean"][rows]) and np.array_equal(ctx.get_A(), D["A_neohookean"][idx])
            ctx.set_r(None)
            ctx.assemble_residual_device()
            ok = ok and np.array_equal(ctx.get_r()[rows], D["r_neohookean"][rows])
            ok_all = ok_all and ok
        # the setup forms ride the same pipelines and the same ghost exchange: mass / viscosity (K_e blocks), body force (R_e)
        from tests.test_gpu_mass import bodyforce, patch_of
        G = bodyforce()
        ncp, knots, ijk = patch_of(D, 2)
        mine = ijk[loc[:R["n_compute"]]]
        ctx.set_nodes(D["x_ref"], None)
        ctx.set_A(None)
        ctx.assemble_mass_device(2, ncp, knots, mine)
        ok = np.array_equal(ctx.get_A(), D["mass"][idx])
        ctx.set_A(None)
        ctx.assemble_viscosity_device(2, ncp, knots, mine, float(G["viscosity"]))
        ok = ok and np.array_equal(ctx.get_A(), G["visc_" + name][idx])
        ctx.set_r(None)
        ctx.assemble_body_force_device(2, ncp, knots, mine, G["f"])
        ok = ok and np.array_equal(ctx.get_r()[rows], G[name][rows])
        ok_all = ok_all and ok
        print("rank %d/%d %s forms (mass, viscosity, body force): %s" % (rank, world, name, "OK" if ok else "MISMATCH"), flush=True)
        print("rank %d/%d %s: %s (own %d elements, %d ghosts, %d owned rows)" % (rank, world, name, "OK" if ok_all else "MISMATCH",
              R["n_compute"], len(loc) - R["n_compute"], int(rows.sum())), flush=True)
        ctx.close()
    t = torch.tensor([int(ok_all)], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    sys.exit(0 if int(t) == 1 else 1)


if __name__ == "__main__":
    main()
