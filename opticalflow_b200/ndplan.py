"""Nested-dissection plan for the sparse direct solve of one frame pair (host side, numpy only).

The reference's ``use_direct_solver=True`` branch hands the assembled matrix to SuperLU (optical_flow.py:1146-1147).
On the device the same system - with the ghost ring eliminated, i.e. over the interior pixels - is factorised by a
multifrontal method on a geometric nested dissection of the pixel grid: a rectangle of pixels is cut by a one-pixel
line (its *separator*: the 9-point stencil does not reach across it), the halves are cut again, down to leaves of a
few pixels.  A tree node owns its separator (a leaf: all its pixels); its *border* is the ring of pixels around its
rectangle, which belong to separators higher up.  Eliminating the owned pixels of a node needs a dense *front* over
owned + border pixels (three unknowns per pixel) and leaves a Schur complement on the border that is added into the
parent's front.

The tree is complete and all nodes of a level are padded to the level's largest owned / border counts, so every level is
one batch of equally sized dense problems for the device.  This module only builds the index arrays (the symbolic
phase); it depends on the frame shape alone and is cached per shape.  The numeric phase is ``ofb_direct_*`` in
libofb200.so.
"""
import functools

import numpy as np

MIN_LEAF = 2            # a region is split while both halves keep at least this many pixels along the cut axis


class Level:
    """All nodes of one tree depth.  Index arrays are int32, -1 = padding."""

    def __init__(self, n_nodes, n_own, n_border):
        self.n_nodes, self.n_own, self.n_border = n_nodes, n_own, n_border
        self.own = np.full((n_nodes, n_own), -1, np.int32)          # interior pixel ids (i * my + j) the node eliminates
        self.border = np.full((n_nodes, n_border), -1, np.int32)    # pixel ids of the ring around its rectangle
        self.parent_slot = np.full((n_nodes, n_border), -1, np.int32)   # border pixel p -> position in the PARENT's front
        self.rect = np.zeros((n_nodes, 4), np.int32)                # i0, i1, j0, j1 (half-open)

    @property
    def n_front(self):
        return self.n_own + self.n_border


def _ring(i0, i1, j0, j1, mx, my):
    """Pixel ids of the one-pixel ring around the rectangle, clipped to the grid."""
    ii = np.arange(max(i0 - 1, 0), min(i1 + 1, mx))
    jj = np.arange(max(j0 - 1, 0), min(j1 + 1, my))
    I, J = np.meshgrid(ii, jj, indexing='ij')
    outside = (I < i0) | (I >= i1) | (J < j0) | (J >= j1)
    return (I[outside] * my + J[outside]).astype(np.int64)


def _dissect(mx, my, max_depth):
    """The rectangles of every tree depth and the separator lines that cut them (plain Python: ~2 x 10^5 tuples at 1024^2)."""
    rects = [[(0, mx, 0, my)]]
    seps = []
    while max_depth is None or len(seps) < max_depth:
        # complete tree: a level is cut only if EVERY region of it can be cut with both halves >= MIN_LEAF
        if min(max(i1 - i0, j1 - j0) for (i0, i1, j0, j1) in rects[-1]) < 2 * MIN_LEAF + 1 or len(rects[-1]) >= (1 << 20):
            break
        nxt, sep = [], []
        for (i0, i1, j0, j1) in rects[-1]:
            if (i1 - i0) >= (j1 - j0):
                m = (i0 + i1) // 2
                sep.append((m, m + 1, j0, j1))
                nxt += [(i0, m, j0, j1), (m + 1, i1, j0, j1)]
            else:
                m = (j0 + j1) // 2
                sep.append((i0, i1, m, m + 1))
                nxt += [(i0, i1, j0, m), (i0, i1, m + 1, j1)]
        seps.append(sep)
        rects.append(nxt)
    return rects, seps


def _pack_rows(node, values, n_nodes, floor=1):
    """(node index, value) pairs in node-major order -> [n_nodes, max count] int32, -1 padded, order kept."""
    counts = np.bincount(node, minlength=n_nodes)
    width = max(floor, int(counts.max()) if counts.size else 0)
    out = np.full((n_nodes, width), -1, np.int32)
    start = np.cumsum(counts) - counts
    out[node, np.arange(node.size) - start[node]] = values
    return out


@functools.lru_cache(maxsize=8)
def build(Nx, Ny, max_depth=None):
    """Plan for an Nx x Ny frame (ghost ring included).  Returns a list of ``Level`` from the root (index 0) to the leaves.
    Array operations over all nodes of a level (1 s at 1024^2); ``build_loops`` is the node-by-node statement of the same plan."""
    mx, my = Nx - 2, Ny - 2
    if mx < 1 or my < 1:
        raise ValueError('frame too small')
    rects, seps = _dissect(mx, my, max_depth)
    depth = len(seps)
    M = mx * my
    levels = []
    for d in range(depth + 1):
        R = np.asarray(rects[d], np.int64).reshape(-1, 4)
        O = np.asarray(seps[d] if d < depth else rects[d], np.int64).reshape(-1, 4)        # leaves own their whole rectangle
        n = R.shape[0]
        # owned pixels, row-major over the owned rectangle (a separator is a line, a leaf a few pixels each way)
        h, w = O[:, 1] - O[:, 0], O[:, 3] - O[:, 2]
        di = np.arange(int(h.max()))[None, :, None]
        dj = np.arange(int(w.max()))[None, None, :]
        ok = (di < h[:, None, None]) & (dj < w[:, None, None])
        pix = (O[:, 0][:, None, None] + di) * my + (O[:, 2][:, None, None] + dj)
        node, a, b = np.nonzero(ok)
        own = _pack_rows(node, pix[node, a, b], n)
        # the ring around the rectangle in row-major order: the row above, (left, right) of every row, the row below; clipped to the grid
        i0, i1, j0, j1 = R[:, 0][:, None], R[:, 1][:, None], R[:, 2][:, None], R[:, 3][:, None]
        hh, ww = i1 - i0, j1 - j0
        t1, t2 = ww + 2, ww + 2 + 2 * hh
        k = np.arange(int((2 * (ww + 2) + 2 * hh).max()))[None, :]
        top, mid, bot = k < t1, (k >= t1) & (k < t2), (k >= t2) & (k < t2 + ww + 2)
        r, side = (k - t1) // 2, (k - t1) % 2
        I = np.where(top, i0 - 1, np.where(mid, i0 + r, i1))
        J = np.where(top, j0 - 1 + k, np.where(mid, np.where(side == 0, j0 - 1, j1), j0 - 1 + (k - t2)))
        ok = (top | mid | bot) & (I >= 0) & (I < mx) & (J >= 0) & (J < my)
        node, pos = np.nonzero(ok)
        border = _pack_rows(node, (I * my + J)[node, pos], n)
        lv = Level(n, own.shape[1], border.shape[1])
        lv.own, lv.border = own, border
        lv.parent_slot = np.full(border.shape, -1, np.int32)
        lv.rect = R.astype(np.int32)
        levels.append(lv)
    # where a node's border pixels sit in its parent's front (own pixels first, then the parent's border; padding slots count)
    for d in range(1, depth + 1):
        lv, par = levels[d], levels[d - 1]
        front = np.concatenate([par.own, par.border], axis=1).astype(np.int64)                 # [parents, front slots]
        p_idx, slot = np.nonzero(front >= 0)
        keys = p_idx * M + front[p_idx, slot]                                                  # (parent, pixel): unique
        order = np.argsort(keys, kind='stable')
        keys, slot = keys[order], slot[order]
        c_idx, c_pos = np.nonzero(lv.border >= 0)
        want = (c_idx // 2) * M + lv.border[c_idx, c_pos].astype(np.int64)
        at = np.minimum(np.searchsorted(keys, want), keys.size - 1)
        if not np.array_equal(keys[at], want):
            raise AssertionError('border pixel of a node is not in its parent front')
        lv.parent_slot[c_idx, c_pos] = slot[at]
    return levels


def build_loops(Nx, Ny, max_depth=None):
    """The same plan node by node (the statement ``build`` is checked against in tests/test_ndplan_cpu.py; 10 s at 1024^2)."""
    mx, my = Nx - 2, Ny - 2
    if mx < 1 or my < 1:
        raise ValueError('frame too small')
    rects, seps = _dissect(mx, my, max_depth)
    depth = len(seps)
    levels = []
    for d in range(depth + 1):
        owned = seps[d] if d < depth else rects[d]            # leaves own their whole rectangle
        own_lists, border_lists = [], []
        for (i0, i1, j0, j1), (a0, a1, b0, b1) in zip(rects[d], owned):
            I, J = np.meshgrid(np.arange(a0, a1), np.arange(b0, b1), indexing='ij')
            own_lists.append((I * my + J).ravel().astype(np.int64))
            border_lists.append(_ring(i0, i1, j0, j1, mx, my))
        lv = Level(len(own_lists), max(len(o) for o in own_lists), max(1, max(len(b) for b in border_lists)))
        for n, (o, b) in enumerate(zip(own_lists, border_lists)):
            lv.own[n, :len(o)] = o
            lv.border[n, :len(b)] = b
            lv.rect[n] = rects[d][n]
        levels.append(lv)
    # where a node's border pixels sit in its parent's front (own pixels first, then the parent's border)
    for d in range(1, depth + 1):
        lv, par = levels[d], levels[d - 1]
        for n in range(lv.n_nodes):
            p = n // 2
            front = np.concatenate([par.own[p], par.border[p]])
            order = np.argsort(front, kind='stable')
            b = lv.border[n]
            valid = b >= 0
            pos = np.searchsorted(front[order], b[valid])
            slot = order[np.minimum(pos, len(order) - 1)]
            if not np.array_equal(front[slot], b[valid]):
                raise AssertionError('border pixel of a node is not in its parent front')
            lv.parent_slot[n, valid] = slot
    return levels


def memory_bytes(levels, itemsize=8):
    """Bytes of device memory the fronts of all levels take (three unknowns per pixel)."""
    return int(sum(lv.n_nodes * (3 * lv.n_front) ** 2 * itemsize for lv in levels))
