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
