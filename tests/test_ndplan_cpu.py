"""The nested-dissection plan of the sparse direct solve (opticalflow_b200/ndplan.py): host-side bookkeeping, CPU only."""
import numpy as np
import pytest

from opticalflow_b200 import ndplan


@pytest.mark.parametrize('shape', [(5, 5), (7, 6), (23, 31), (40, 56), (64, 64), (33, 130), (129, 257), (256, 256)])
def test_plan_partitions_the_interior_and_nests_the_borders(shape):
    Nx, Ny = shape
    mx, my = Nx - 2, Ny - 2
    levels = ndplan.build(Nx, Ny)
    owned = np.concatenate([lv.own[lv.own >= 0] for lv in levels])
    assert np.array_equal(np.sort(owned), np.arange(mx * my))                       # every interior pixel eliminated exactly once
    for d, lv in enumerate(levels):
        assert lv.n_nodes == 1 << d and lv.own.shape == (lv.n_nodes, lv.n_own) and lv.border.shape == (lv.n_nodes, lv.n_border)
        for n in range(lv.n_nodes):
            own, border = lv.own[n][lv.own[n] >= 0], lv.border[n][lv.border[n] >= 0]
            assert len(own) >= 1 and len(np.unique(border)) == len(border) and not np.intersect1d(own, border).size
            # the stencil of an owned pixel (3 x 3 neighbourhood) stays inside the node's rectangle + ring
            i0, i1, j0, j1 = lv.rect[n]
            oi, oj = own // my, own % my
            assert (oi >= i0).all() and (oi < i1).all() and (oj >= j0).all() and (oj < j1).all()
            bi, bj = border // my, border % my
            assert ((bi >= i0 - 1) & (bi <= i1) & (bj >= j0 - 1) & (bj <= j1)).all()
            if d > 0:                                                               # a border pixel sits in the parent's front
                par = levels[d - 1]
                front = np.concatenate([par.own[n // 2], par.border[n // 2]])
                slot = lv.parent_slot[n][lv.border[n] >= 0]
                assert (slot >= 0).all() and np.array_equal(front[slot], border)
    assert levels[0].border.max() == -1                                             # the root has nothing above it


def test_plan_sizes_at_the_bench_shapes():
    levels = ndplan.build(256, 256)
    assert len(levels) == 13 and (levels[0].n_own, levels[-1].n_nodes) == (254, 4096)
    assert 1.0e9 < ndplan.memory_bytes(levels) < 1.5e9
    small = ndplan.build(5, 5)
    assert len(small) == 1 and small[0].n_own == 9                                 # 3 x 3 interior: one dense front


@pytest.mark.parametrize('shape', [(5, 5), (7, 6), (6, 200), (23, 31), (33, 130), (300, 44), (129, 257), (256, 256)])
def test_array_builder_equals_the_node_by_node_statement(shape):
    """``build`` works on all nodes of a level at once; ``build_loops`` states the same plan node by node: same arrays, same order."""
    fast, slow = ndplan.build(*shape), ndplan.build_loops(*shape)
    assert len(fast) == len(slow)
    for a, b in zip(fast, slow):
        assert (a.n_nodes, a.n_own, a.n_border) == (b.n_nodes, b.n_own, b.n_border)
        for name in ('own', 'border', 'parent_slot', 'rect'):
            x, y = getattr(a, name), getattr(b, name)
            assert x.dtype == y.dtype and np.array_equal(x, y), name
