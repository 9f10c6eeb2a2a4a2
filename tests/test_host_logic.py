"""CPU: host-side mirror of the reference interface (argument checks, settings tuple, view sharding, scenes)."""
import numpy as np
import pytest
import torch


def test_settings_fields_match_reference_order():
    # DGR/diff_gaussian_rasterization_rade/__init__.py:177-192
    from igs_b200 import GaussianRasterizationSettings as S
    assert S._fields == ("image_height", "image_width", "tanfovx", "tanfovy", "kernel_size", "bg", "scale_modifier",
                         "viewmatrix", "projmatrix", "sh_degree", "campos", "prefiltered", "require_depth",
                         "require_coord", "debug")


def _settings():
    from igs_b200 import GaussianRasterizationSettings as S
    return S(16, 16, 1.0, 1.0, 0.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 3, torch.zeros(3), False, True, True,
             False)


def test_exactly_one_of_checks():
    # reference __init__.py:214-218: plain Exception on argument-pair misuse
    from igs_b200 import GaussianRasterizer
    r = GaussianRasterizer(_settings())
    m = torch.zeros(2, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=m, means2D=m, opacities=torch.ones(2, 1), scales=torch.ones(2, 3), rotations=torch.ones(2, 4))
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(means3D=m, means2D=m, opacities=torch.ones(2, 1), shs=torch.zeros(2, 16, 3), colors_precomp=torch.zeros(2, 3),
          scales=torch.ones(2, 3), rotations=torch.ones(2, 4))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m, opacities=torch.ones(2, 1), shs=torch.zeros(2, 16, 3))
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m, means2D=m, opacities=torch.ones(2, 1), shs=torch.zeros(2, 16, 3), scales=torch.ones(2, 3),
          rotations=torch.ones(2, 4), cov3D_precomp=torch.zeros(2, 6))


def test_integrate_is_out_of_scope():
    from igs_b200 import GaussianRasterizer
    with pytest.raises(NotImplementedError):
        GaussianRasterizer(_settings()).integrate()


def test_drop_in_packages_expose_reference_names():
    import diff_gaussian_rasterization_rade as a
    import diff_gaussian_rasterization_rade_clamp as b
    for mod in (a, b):
        for n in ("GaussianRasterizationSettings", "GaussianRasterizer", "_RasterizeGaussians", "rasterize_gaussians", "_C"):
            assert hasattr(mod, n), (mod.__name__, n)
    assert b._RasterizeGaussians.clamp_grads and not a._RasterizeGaussians.clamp_grads


def test_shard_views():
    from igs_b200.refine import shard_views
    counts = [len(shard_views(20, r, 8)) for r in range(8)]
    assert counts == [3, 3, 3, 3, 2, 2, 2, 2]
    allv = sorted(v for r in range(8) for v in shard_views(20, r, 8))
    assert allv == list(range(20))
    assert shard_views(3, 5, 8) == []


def test_shard_views_balanced():
    from igs_b200.refine import shard_views_balanced
    costs = [10, 9, 8, 7, 6, 5, 4, 3, 2, 1]
    parts = [shard_views_balanced(costs, r, 3) for r in range(3)]
    assert sorted(v for p in parts for v in p) == list(range(10))
    loads = [sum(costs[v] for v in p) for p in parts]
    assert max(loads) - min(loads) <= 2
    assert shard_views_balanced([5, 5], 2, 3) == []


def test_scenes_are_deterministic_and_cameras_follow_reference_convention():
    from igs_b200 import scenes
    a, b = scenes.make_gaussians(100, 0), scenes.make_gaussians(100, 0)
    assert torch.equal(a.xyz, b.xyz) and torch.equal(a.sh, b.sh)
    cams = scenes.make_cameras(3, 64, 48, 1)
    # camera centre = inverse(world_view)[3,:3] (igs/models/gs.py:106); looking at LOOK_AT gives z_view > 0
    for v in range(3):
        p = torch.tensor([*scenes.LOOK_AT, 1.0]) @ cams.viewmatrix[v]
        assert p[2] > 0 and abs(p[0]) < 1e-4 and abs(p[1]) < 1e-4
        o = torch.cat([cams.campos[v], torch.ones(1)]) @ cams.viewmatrix[v]
        assert o[:3].abs().max() < 1e-4
    assert abs(cams.tanfovy - cams.tanfovx * 48 / 64) < 1e-6


def test_refine_packing_constants():
    from igs_b200 import refine
    assert refine.FLOATS_PER_GAUSSIAN == 59  # xyz 3 + rot 4 + sh 48 + opacity 1 + scale 3
