"""CPU, world_size 2 over gloo: the view-sharded exchange step (igs_b200.refine.reduce_gradients) gives every
rank the same packed gradient as a single process summing all views.  The per-view gradients come from the CPU
oracle here (the CUDA path needs a GPU); the N-rank == 1-rank property on the real kernels is in
tests/test_refine_gpu.py."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent
V, P, W, H = 4, 300, 48, 32


def _view_grads(v):
    for p in (ROOT, ROOT / "tests", ROOT / "oracle"):
        if str(p) not in sys.path:
            sys.path.insert(0, str(p))
    import oracle_py
    from igs_b200 import scenes
    sc = scenes.make_gaussians(P, 5, -1.2)
    cams = scenes.make_cameras(V, W, H, 6)
    scales, rot, opac = scenes.activated(sc)
    fw = oracle_py.forward(sc.xyz.numpy(), scales.numpy(), rot.numpy(), opac.numpy(), sc.sh.numpy(), cams.viewmatrix[v].numpy(),
                           cams.projmatrix[v].numpy(), cams.campos[v].numpy(), W, H, cams.tanfovx, cams.tanfovy)
    rng = np.random.RandomState(100 + v)
    g = oracle_py.backward(fw, dL_color=rng.standard_normal((3, H, W)).astype(np.float32))
    loss = float(fw["image"]["color"].sum())
    return g, loss


def _packed(views):
    from igs_b200 import refine
    flat = torch.zeros(P * refine.FLOATS_PER_GAUSSIAN)
    loss = torch.zeros(())
    for v in views:
        g, l = _view_grads(v)
        flat += refine.pack_gradients(dict(xyz=torch.from_numpy(g["means3D"]), rotation=torch.from_numpy(g["rotations"]),
                                           shs=torch.from_numpy(np.ascontiguousarray(g["sh"])),
                                           opacity=torch.from_numpy(g["opacities"]), scaling=torch.from_numpy(g["scales"])), P)
        loss += l / V
    return flat, loss


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from igs_b200 import refine
    flat, loss = _packed(refine.shard_views(V, rank, world))
    loss = refine.reduce_gradients(flat, loss, dist.group.WORLD)
    torch.save((flat, loss), f"{out}/r{rank}.pt")
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_reduction_equals_single_process(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    f0, l0 = torch.load(tmp_path / "r0.pt")
    f1, l1 = torch.load(tmp_path / "r1.pt")
    assert torch.equal(f0, f1) and torch.equal(l0, l1)  # identical inputs to the (identical) Adam step on every rank
    fs, ls = _packed(range(V))
    assert float((f0 - fs).abs().max()) <= 1e-5 * float(fs.abs().max())
    assert abs(float(l0) - float(ls)) <= 1e-5 * abs(float(ls))
    assert float(fs.abs().max()) > 0


def _densify_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from igs_b200 import densify, refine
    g = torch.Generator().manual_seed(9)
    Pn, Vn = 900, 4
    g2d = torch.randn(Vn, Pn, 3, generator=g) * 2e-4
    radii = torch.randint(-1, 5, (Vn, Pn), generator=g, dtype=torch.int32).clamp_min(0)
    log_scale = torch.randn(Pn, 3, generator=g) * 0.8 - 3.0
    logit = torch.randn(Pn, 1, generator=g) * 3
    views = refine.shard_views(Vn, rank, world)
    st = densify.DensifyStats(Pn, "cpu")
    st.update_torch(g2d[views], radii[views])
    st.all_reduce(dist.group.WORLD)
    plan = densify.plan_densify(None, log_scale, logit, st.grad_accum, st.denom, 0.00015, 0.005, 3.0, None, 150000, 0.01,
                                generator=torch.Generator().manual_seed(77))
    torch.save((st.grad_accum, st.denom, st.max_radii2D, plan.src, plan.kind, plan.samples), f"{out}/d{rank}.pt")
    if rank == 0:
        full = densify.DensifyStats(Pn, "cpu")
        full.update_torch(g2d, radii)
        torch.save((full.grad_accum, full.denom, full.max_radii2D), f"{out}/dfull.pt")
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_densification_plans_are_identical(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_densify_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a, b = torch.load(tmp_path / "d0.pt"), torch.load(tmp_path / "d1.pt")
    for x, y in zip(a, b):
        assert torch.equal(x, y)                      # same statistics -> same plan on every rank, no broadcast
    full = torch.load(tmp_path / "dfull.pt")
    assert torch.allclose(a[0], full[0], rtol=1e-6, atol=1e-9) and torch.equal(a[1], full[1]) and torch.equal(a[2], full[2])
    assert int((a[4] == 1).sum()) > 0 and int((a[4] == 2).sum()) > 0
