"""CPU: host logic of the key-frame streaming loop (igs_b200/stream.py) against a restatement of the reference's
loop bookkeeping (infer_batch.py:279-324)."""
import random

import pytest

torch = pytest.importorskip("torch")


def _stream():
    try:
        from igs_b200 import stream
    except ImportError as ex:  # the package needs its built extension even on CPU
        pytest.skip(f"igs_b200 not importable: {ex}")
    return stream


def _reference_draws(n_views, iterations, seed):
    """infer_batch.py:283-292 verbatim in spirit: copy the list when empty, randint + pop."""
    from random import randint, seed as rseed
    rseed(seed)
    cams, out = None, []
    for _ in range(iterations):
        if not cams:
            cams = list(range(n_views))
        pick = randint(0, len(cams) - 1)
        out.append(cams.pop(pick))
    return out


@pytest.mark.parametrize("V,iters", [(13, 50), (20, 50), (1, 7), (3, 2)])
def test_view_schedule_is_the_references_draw(V, iters):
    s = _stream()
    got = s.view_schedule(V, iters, random.Random(123))
    assert [g[0] for g in got] == _reference_draws(V, iters, 123)
    # without replacement: every full pass over the list visits each view exactly once
    flat = [g[0] for g in got]
    for a in range(0, iters - V + 1, V):
        assert sorted(flat[a:a + V]) == list(range(V))


def test_view_schedule_batched_steps_cover_all_views():
    s = _stream()
    got = s.view_schedule(20, 3, random.Random(0), views_per_step=20)
    assert all(sorted(g) == list(range(20)) for g in got)


def test_select_views_slices_without_copy():
    s = _stream()
    from igs_b200.multiview import MultiViewSettings
    vm = torch.arange(5 * 16, dtype=torch.float32).view(5, 4, 4)
    st = MultiViewSettings(8, 8, 1.0, 1.0, 0.0, torch.zeros(3), 1.0, vm, vm + 1, 3, torch.zeros(5, 3))
    one = s.select_views(st, [3])
    assert one.viewmatrices.shape == (1, 4, 4) and one.viewmatrices.data_ptr() == vm[3:4].data_ptr()
    run = s.select_views(st, [1, 2, 3])
    assert run.viewmatrices.data_ptr() == vm[1:4].data_ptr()
    mix = s.select_views(st, [4, 0])
    assert torch.equal(mix.projmatrices, (vm + 1)[[4, 0]]) and mix.viewmatrices.is_contiguous()


def test_densify_iterations_follow_the_reference_condition():
    """iteration > densify_from_iter and iteration % interval == 0, while iteration < densify_until_iter."""
    s = _stream()
    cfg = s.StreamConfig(use_densify=True)
    hits = [it for it in range(cfg.refine_iterations)
            if it < cfg.densify_until_iter and it > cfg.densify_from_iter and it % cfg.densification_interval == 0]
    assert hits == [20, 40]
