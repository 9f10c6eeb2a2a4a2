"""GPU: size-independent properties at BASELINE.json's full sizes (C2 200k/512^2 single view; one C3 view
300k/1352x1014), where the CPU oracle would be too slow: sortedness, range partition, linearity of backward in
the cotangent, determinism of the integer outputs, bounds."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _scene(cfg_name, dev, view=0):
    from igs_b200 import scenes
    cfg = scenes.CONFIGS[cfg_name]
    sc = scenes.make_gaussians(cfg["P"], 0, cfg["log_scale_mean"])
    cams = scenes.make_cameras(cfg["V"], cfg["W"], cfg["H"], 1)
    scales, rot, opac = scenes.activated(sc)
    t = dict(means3D=sc.xyz.to(dev), shs=sc.sh.to(dev), opacities=opac.to(dev), scales=scales.to(dev), rotations=rot.to(dev))
    cam = dict(vm=cams.viewmatrix[view].to(dev), pm=cams.projmatrix[view].to(dev), cp=cams.campos[view].to(dev),
               tx=cams.tanfovx, ty=cams.tanfovy)
    return cfg, t, cam


@pytest.mark.parametrize("cfg_name,view", [("C2", 0), ("C3", 7), ("C4", 5), ("C5", 3)])
def test_binning_invariants_full_size(cuda_dev, native, cfg_name, view):
    cfg, t, cam = _scene(cfg_name, cuda_dev, view)
    P, W, H = cfg["P"], cfg["W"], cfg["H"]
    e = torch.Tensor([])
    bg = torch.zeros(3, device=cuda_dev)
    nat = native.rasterize_gaussians(bg, t["means3D"], e, t["opacities"], t["scales"], t["rotations"], 1.0, e, cam["vm"], cam["pm"],
                                     cam["tx"], cam["ty"], 0.0, H, W, t["shs"], 3, cam["cp"], False, True, True, False)
    R = nat[0]
    keys, plist, ranges, ncontrib, tiles_t, depths = native.decode_buffers(P, 1, R, H, W, nat[9], nat[10], nat[11])[:6]
    radii = nat[8]
    assert R > 0 and int(tiles_t.sum()) == R
    assert bool(((radii > 0) == (tiles_t.reshape(-1) > 0)).all())
    # sorted ascending on the 64-bit (tile | depth) key; ties in ascending Gaussian index (stable)
    k = keys.to(torch.int64)
    assert bool((k[1:] >= k[:-1]).all())
    tie = k[1:] == k[:-1]
    assert bool((plist[1:][tie] > plist[:-1][tie]).all())
    # depth bits of every key equal the Gaussian's view depth; tiles are inside the grid
    d = depths.reshape(-1)[plist.long()].view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    assert bool(((k & 0xFFFFFFFF) == d).all())
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    tile_of = k >> 32
    assert int(tile_of.max()) < tiles
    # ranges partition [0, R): tile t owns exactly the instances whose key's tile is t
    r = ranges.to(torch.int64)
    counts = torch.bincount(tile_of, minlength=tiles)
    assert bool(((r[:, 1] - r[:, 0]) == counts).all())
    nz = counts > 0
    assert bool((tile_of[r[nz, 0]] == torch.nonzero(nz).reshape(-1)).all())
    # n_contrib: last contributor within the tile's list length; outputs bounded
    nc = ncontrib.reshape(2, H, W)[0].to(torch.int64)
    per_pix_len = counts.reshape((H + 15) // 16, (W + 15) // 16).repeat_interleave(16, 0).repeat_interleave(16, 1)[:H, :W]
    assert bool((nc <= per_pix_len).all())
    alpha, color = nat[4], nat[1]
    assert float(alpha.min()) >= 0 and float(alpha.max()) <= 1.0 + 1e-5
    assert bool(torch.isfinite(color).all())
    # determinism of everything that must be bit-exact
    nat2 = native.rasterize_gaussians(bg, t["means3D"], e, t["opacities"], t["scales"], t["rotations"], 1.0, e, cam["vm"], cam["pm"],
                                      cam["tx"], cam["ty"], 0.0, H, W, t["shs"], 3, cam["cp"], False, True, True, False)
    assert nat2[0] == R
    for i in range(1, 9):
        assert torch.equal(nat[i], nat2[i]), i


def test_backward_is_linear_in_the_cotangent(cuda_dev, native):
    import igs_b200
    cfg, t, cam = _scene("C2", cuda_dev)
    W, H = cfg["W"], cfg["H"]
    bg = torch.tensor([0.2, 0.4, 0.6], device=cuda_dev)
    st = igs_b200.GaussianRasterizationSettings(H, W, cam["tx"], cam["ty"], 0.0, bg, 1.0, cam["vm"], cam["pm"], 3, cam["cp"], False,
                                                True, True, False)
    g = torch.Generator(device="cpu").manual_seed(11)
    c1 = torch.randn(3, H, W, generator=g).to(cuda_dev)
    c2 = torch.randn(3, H, W, generator=g).to(cuda_dev)

    def grads(cot):
        leaves = {k: v.clone().requires_grad_(True) for k, v in t.items()}
        m2 = torch.zeros_like(leaves["means3D"], requires_grad=True)
        out = igs_b200.GaussianRasterizer(st)(means2D=m2, **leaves)
        (out[0] * cot).sum().backward()
        return {k: v.grad for k, v in leaves.items()}

    ga, gb, gc = grads(c1), grads(c2), grads(2.0 * c1 - 0.5 * c2)
    for k in ga:
        lin = 2.0 * ga[k] - 0.5 * gb[k]
        den = float(gc[k].abs().max()) + 1e-30
        assert float((lin - gc[k]).abs().max()) / den <= 1e-3, k


def test_zero_cotangent_gives_zero_gradients_and_culled_gaussians_get_none(cuda_dev, native):
    import igs_b200
    cfg, t, cam = _scene("C1", cuda_dev)
    W, H = cfg["W"], cfg["H"]
    st = igs_b200.GaussianRasterizationSettings(H, W, cam["tx"], cam["ty"], 0.0, torch.zeros(3, device=cuda_dev), 1.0, cam["vm"],
                                                cam["pm"], 3, cam["cp"], False, True, True, False)
    leaves = {k: v.clone().requires_grad_(True) for k, v in t.items()}
    m2 = torch.zeros_like(leaves["means3D"], requires_grad=True)
    out = igs_b200.GaussianRasterizer(st)(means2D=m2, **leaves)
    (out[0] * 0.0).sum().backward()
    for k, v in leaves.items():
        assert float(v.grad.abs().max()) == 0.0, k
    leaves = {k: v.clone().requires_grad_(True) for k, v in t.items()}
    out = igs_b200.GaussianRasterizer(st)(means2D=m2, **leaves)
    out[0].sum().backward()
    culled = out[1] == 0
    assert int(culled.sum()) > 0
    for k, v in leaves.items():
        assert float(v.grad[culled].abs().max()) == 0.0, k


def test_empty_and_tiny_inputs(cuda_dev, native):
    import igs_b200
    cfg, t, cam = _scene("C1", cuda_dev)
    W, H = 40, 24
    st = igs_b200.GaussianRasterizationSettings(H, W, cam["tx"], cam["ty"], 0.0, torch.ones(3, device=cuda_dev), 1.0, cam["vm"],
                                                cam["pm"], 3, cam["cp"], False, True, True, False)
    # P == 0: outputs zero, rendered 0 (rasterize_points.cu:89-91)
    z = {k: v[:0].clone().requires_grad_(True) for k, v in t.items()}
    out = igs_b200.GaussianRasterizer(st)(means2D=torch.zeros(0, 3, device=cuda_dev), **z)
    assert out[0].shape == (3, H, W) and float(out[0].abs().max()) == 0.0 and out[1].numel() == 0
    # P == 1
    one = {k: v[:1].clone().requires_grad_(True) for k, v in t.items()}
    out = igs_b200.GaussianRasterizer(st)(means2D=torch.zeros(1, 3, device=cuda_dev), **one)
    out[0].sum().backward()
    assert bool(torch.isfinite(out[0]).all())
    # markVisible == (view z > 0.2)
    vis = igs_b200.GaussianRasterizer(st).markVisible(t["means3D"])
    z_view = (torch.cat([t["means3D"], torch.ones_like(t["means3D"][:, :1])], 1) @ cam["vm"])[:, 2]
    sure = (z_view - 0.2).abs() > 1e-4
    assert bool((vis[sure] == (z_view[sure] > 0.2)).all())
    # prefiltered=True is rejected loudly
    st2 = st._replace(prefiltered=True)
    with pytest.raises(RuntimeError, match="prefiltered"):
        igs_b200.GaussianRasterizer(st2)(means2D=torch.zeros(1, 3, device=cuda_dev), **one)
