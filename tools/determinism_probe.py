"""One run of a full-size case saved to disk (`run <case> <file>`), or a bit-exact comparison of two such files
(`cmp <a> <b>`): see tools/determinism_probe.sh.  Found that torch's float32 CPU exp made the TEST INPUTS differ in the
last bit between processes (tests/fullsize.py:_activated_reproducible); the library itself is bit-reproducible."""
import sys, torch
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import fullsize
dev = torch.device("cuda:0")
if sys.argv[1] == "run":
    inp = fullsize.inputs(sys.argv[2], dev)
    from igs_b200 import _C as native
    out = fullsize.run_ours(inp, native)
    torch.cuda.synchronize()
    cur = {"int_" + k: [t.cpu() for t in v] for k, v in out["ints"].items()}
    cur.update({"map_" + k: v.cpu() for k, v in out["maps"].items()})
    cur.update({"grad_" + k: v.cpu() for k, v in out["grads"].items()})
    cur.update({"in_" + k: v.cpu() for k, v in inp.items() if torch.is_tensor(v)})
    for k, v in inp.items():
        if isinstance(v, (list, tuple)) and len(v) and torch.is_tensor(v[0]):
            cur["in_" + k] = [t.cpu() for t in v]
    torch.save(cur, sys.argv[3])
else:
    a = torch.load(sys.argv[2]); b = torch.load(sys.argv[3])
    bad = []
    for k in a:
        if isinstance(a[k], list): same = all(torch.equal(x, y) for x, y in zip(a[k], b[k]))
        else: same = torch.equal(a[k], b[k])
        if not same: bad.append(k)
    print(sys.argv[2], "vs", sys.argv[3], "differ in:", bad)
    for k in bad:
        x, y = (a[k][0], b[k][0]) if isinstance(a[k], list) else (a[k], b[k])
        d = (x.double() - y.double()).abs()
        idx = torch.nonzero(d.flatten() > 0).flatten()
        print("   ", k, tuple(x.shape), "n differing elems", idx.numel(), "max abs", float(d.max()), "first flat idx", idx[:8].tolist())
