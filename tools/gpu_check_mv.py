"""GPU check: the batched V-view call equals V single-view calls (maps bit-exact, grads summed)."""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from igs_b200 import scenes  # noqa: E402
import igs_b200  # noqa: E402
from igs_b200.multiview import MultiViewSettings, rasterize_views  # noqa: E402


def main():
    P, W, H, V = 20000, 200, 152, 3
    dev = torch.device("cuda:0")
    sc = scenes.make_gaussians(P, 0, -2.0)
    cams = scenes.make_cameras(V, W, H, 1)
    scales, rot, opac = scenes.activated(sc)
    base = [sc.xyz.to(dev), sc.sh.to(dev), opac.to(dev), scales.to(dev), rot.to(dev)]
    bg = torch.zeros(3, device=dev)
    cot = scenes.make_cotangents(V, H, W, 3)[0].to(dev)

    leaves = [b.clone().requires_grad_(True) for b in base]
    st = MultiViewSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, bg, 1.0, cams.viewmatrix.to(dev), cams.projmatrix.to(dev),
                           3, cams.campos.to(dev))
    outs = rasterize_views(*leaves, st)
    (outs[0] * cot).sum().backward()
    g_mv = [l.grad.clone() for l in leaves]

    leaves2 = [b.clone().requires_grad_(True) for b in base]
    maps = []
    for v in range(V):
        s1 = igs_b200.GaussianRasterizationSettings(H, W, cams.tanfovx, cams.tanfovy, 0.0, bg, 1.0,
                                                    cams.viewmatrix[v].to(dev), cams.projmatrix[v].to(dev), 3,
                                                    cams.campos[v].to(dev), False, True, True, False)
        m2 = torch.zeros(P, 3, device=dev, requires_grad=True)
        o = igs_b200.GaussianRasterizer(s1)(means3D=leaves2[0], means2D=m2, shs=leaves2[1], opacities=leaves2[2],
                                            scales=leaves2[3], rotations=leaves2[4])
        maps.append(o)
        (o[0] * cot[v]).sum().backward()
    ok = True
    for i, n in enumerate(["color", "radii", "coord", "mcoord", "depth", "mdepth", "alpha", "normal"]):
        a = outs[i]
        b = torch.stack([m[i] for m in maps])
        neq = int((a != b).sum())
        print(f"map {n:7s} mismatching elements {neq}")
        ok &= neq == 0
    for n, a, l in zip(["xyz", "sh", "opac", "scales", "rot"], g_mv, leaves2):
        d = float((a - l.grad).abs().max())
        r = d / (float(l.grad.abs().max()) + 1e-30)
        print(f"grad {n:7s} max-abs {d:.3e} rel {r:.3e}")
        ok &= r < 1e-4
    print("MULTIVIEW", "OK" if ok else "MISMATCH")


if __name__ == "__main__":
    main()
