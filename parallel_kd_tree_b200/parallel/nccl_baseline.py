"""NCCL + library-sort baseline (BASELINE.json "the baseline we must beat").

The same tree as the product builds, produced the conventional way:

* distributed levels: radix select whose histograms are summed with ``all_reduce``, points moved with
  ``all_to_all_single``, result gathered with ``all_gather_into_tensor`` — one collective call per step,
  nothing fused;
* every level below: library sorts over the whole local array, keyed by (unit, key) with a stable pre-sort on
  the point index (two ``torch.sort`` calls per level: (segment, key, index) is 88 bits wide and ``torch.sort``
  takes one 64-bit key; no host synchronisation inside the level loop).  A *unit* is what must stay together: the unpinned part of a range, a
  range's pinned last element, or an already placed node; units are numbered by position, so one global
  sort per level sorts every range's first n-1 elements in place (exactly kdtree_sequential.cpp:46-48).

Pure PyTorch: runs on CUDA tensors with the nccl backend and on CPU tensors with gloo (that is how the
logic is tested without GPUs).  This module is a baseline and an independent cross-check; none of the
library's kernels are used here.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..models.forest import slice_bounds

_SIGN = -(1 << 63)


def ordered_key(x: torch.Tensor) -> torch.Tensor:
    """float32 -> int64 in [0, 2^32) with the order of the floats (-0 == +0), like kd::ordered_key."""
    u = x.contiguous().view(torch.int32).to(torch.int64) & 0xFFFFFFFF
    u = torch.where((u & 0x7FFFFFFF) == 0, torch.zeros_like(u), u)
    return torch.where(u >= (1 << 31), u ^ 0xFFFFFFFF, u | (1 << 31))


def composite(key: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """(key, idx) as one int64 whose signed order is the lexicographic order."""
    return ((key << 32) | idx.to(torch.int64)) + _SIGN


def _child(n: int):
    return n // 2, n - n // 2 - 1


def seg_range(n_total: int, level: int, s: int):
    lo, n = 0, n_total
    for b in range(level - 1, -1, -1):
        m = n // 2
        if (s >> b) & 1:
            lo, n = lo + m + 1, n - m - 1
        else:
            n = m
    return lo, n


def sort_build_local(pts: torch.Tensor, idx: torch.Tensor, depth0: int) -> torch.Tensor:
    """Sub-tree of the range held in (pts, idx) (last element pinned) by one library sort per level.

    Returns the order (indices into the inputs) in in-order slot order."""
    n, dim = pts.shape
    dev = pts.device
    order = torch.arange(n, device=dev)
    lo = torch.zeros(n, dtype=torch.int64, device=dev)      # range of every slot: start
    cnt = torch.full((n,), n, dtype=torch.int64, device=dev)  # and size; 0 = the slot holds a placed node
    pos = torch.arange(n, device=dev)
    level = 0
    while (n >> level) >= 3:      # the left-most range is the largest; its size depends on n only (no device read-back)
        axis = (depth0 + level) % dim
        free = (cnt > 0) & (pos < lo + cnt - 1)
        unit = torch.where(free, lo, pos)                    # pinned elements and nodes are their own unit
        key = ordered_key(pts[order, axis])
        key = torch.where(free, key, torch.zeros_like(key))
        by_idx = torch.sort(idx[order], stable=True).indices  # tie-break: point index
        k2 = ((unit << 32) | key)[by_idx]
        srt = torch.sort(k2, stable=True).indices
        order = order[by_idx][srt]
        # descend: slot -> child range
        m = cnt // 2
        node = lo + m
        left = (cnt > 0) & (pos < node)
        right = (cnt > 0) & (pos > node)
        lo = torch.where(right, node + 1, lo)
        cnt = torch.where(left, m, torch.where(right, cnt - m - 1, torch.zeros_like(cnt)))
        level += 1
    return order


def sort_build(points: torch.Tensor) -> torch.Tensor:
    """Single-device library-sort build: perm (int32, in-order layout)."""
    n = points.shape[0]
    idx = torch.arange(n, device=points.device)
    return idx[sort_build_local(points, idx, 0)].to(torch.int32)


def nccl_build(local_pts: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """Distributed build with collectives only; every rank returns the complete perm (int32[n_total])."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    assert world & (world - 1) == 0, "world must be a power of two"
    dev = local_pts.device
    dim = local_pts.shape[1]
    first, local = slice_bounds(n_total, world, rank)
    pts = local_pts
    idx = torch.arange(first, first + local, device=dev)
    perm = torch.full((n_total,), -1, dtype=torch.int64, device=dev)
    levels = world.bit_length() - 1
    for level in range(levels):
        G = world >> level
        gb, j = (rank // G) * G, rank % G
        lo, n = seg_range(n_total, level, rank // G)
        m, nr = _child(n)
        axis = level % dim
        free_n = pts.shape[0] - (1 if j == G - 1 else 0)
        comp = composite(ordered_key(pts[:free_n, axis]), idx[:free_n])
        # ---- radix select of rank m over the group's n-1 unpinned composites (all_reduce'd histograms) ----
        prefix, want, shift = 0, m, 64
        u = comp - _SIGN  # back to [0, 2^64) as two's complement; digits are taken with logical shifts below
        for d, (sh, bits) in enumerate(((53, 11), (42, 11), (32, 10), (21, 11), (10, 11), (0, 10))):
            mask = torch.ones_like(u, dtype=torch.bool) if d == 0 else (_lsr(u, shift) == (prefix >> shift))
            digit = _lsr(u, sh) & ((1 << bits) - 1)
            hist = torch.bincount(digit[mask], minlength=2048)[:2048]
            hist = _group_all_reduce(hist, gb, G, world, group)
            cum = torch.cumsum(hist, 0) - hist
            b = int(((cum <= want) & (want < cum + hist)).nonzero()[0])
            prefix |= b << sh
            want -= int(cum[b])
            shift = sh   # no early exit: every group must issue the same sequence of collectives
        if shift > 0:
            c, p = _lsr(u, shift), prefix >> shift
        else:                       # all 64 bits resolved: compare in the signed-ordered domain
            c, p = comp, prefix + _SIGN
        is_left, is_right, is_node = c < p, c > p, c == p
        # ---- node to every rank's tree array; left maximum identified by an all_reduce(MAX) ----
        node_val = torch.full((1,), -1, dtype=torch.int64, device=dev)
        if bool(is_node.any()):
            node_val[0] = idx[:free_n][is_node][0]
        all_nodes = _group_max(node_val, gb, G, world, group, full=True)   # every group's node, on every rank
        for s_ in range(world // G):
            slo, sn = seg_range(n_total, level, s_)
            perm[slo + sn // 2] = all_nodes[s_, 0]
        lmax = comp[is_left].max().reshape(1) if bool(is_left.any()) else torch.full((1,), _SIGN, dtype=torch.int64, device=dev)
        lmax = _group_max(lmax, gb, G, world, group)
        is_p = is_left & (comp == lmax[0])
        is_left = is_left & ~is_p
        # ---- redistribution with all_to_all_single: left range on the first G/2 members, right on the others ----
        Gh = G // 2
        counts = torch.tensor([int(is_left.sum()), int(is_right.sum())], device=dev)
        allc = _group_all_gather(counts, gb, G, world, group)          # [G, 2]
        Loff, Roff = int(allc[:j, 0].sum()), int(allc[:j, 1].sum())
        # points travel as float32 rows, indices as a separate int64 all_to_all (exact)
        send_rows, send_idx, split = [], [], [0] * world

        def push(sel_pts, sel_idx, off0, n_c, half):
            k = sel_pts.shape[0]
            if k == 0:
                return
            offs = torch.arange(off0, off0 + k, device=dev)
            base, rem = divmod(n_c, Gh)
            big = rem * (base + 1)
            own = torch.where(offs < big, offs // (base + 1), rem + (offs - big) // max(base, 1))
            for t in range(Gh):
                sel = own == t
                cnt_t = int(sel.sum())
                if cnt_t:
                    dst = gb + half * Gh + t
                    send_rows.append((dst, sel_pts[sel]))
                    send_idx.append((dst, sel_idx[sel]))
        fp, fi = pts[:free_n], idx[:free_n]
        push(fp[is_left], fi[is_left], Loff, m, 0)
        if bool(is_p.any()):                                              # the left maximum: last slot of the left range
            send_rows.append((gb + Gh - 1, fp[is_p])); send_idx.append((gb + Gh - 1, fi[is_p]))
        push(fp[is_right], fi[is_right], Roff, nr, 1)
        if j == G - 1:                                                    # the pinned element: last slot of the right range
            send_rows.append((gb + G - 1, pts[free_n:])); send_idx.append((gb + G - 1, idx[free_n:]))
        # assemble per-destination buffers (stable: left chunk, then p, then right chunk, then pinned)
        out_rows, out_idx = [], []
        for dst in range(world):
            rows = [r for d_, r in send_rows if d_ == dst]
            ids = [r for d_, r in send_idx if d_ == dst]
            split[dst] = sum(r.shape[0] for r in rows)
            if rows:
                out_rows.append(torch.cat(rows)); out_idx.append(torch.cat(ids))
        send_p = torch.cat(out_rows) if out_rows else pts[:0]
        send_i = torch.cat(out_idx) if out_idx else idx[:0]
        split_t = torch.tensor(split, device=dev)
        recv_split = torch.empty_like(split_t)
        dist.all_to_all_single(recv_split, split_t, group=group)
        rs = [int(x) for x in recv_split.tolist()]
        recv_p = torch.empty((sum(rs), dim), dtype=pts.dtype, device=dev)
        recv_i = torch.empty((sum(rs),), dtype=torch.int64, device=dev)
        dist.all_to_all_single(recv_p, send_p, output_split_sizes=rs, input_split_sizes=split, group=group)
        dist.all_to_all_single(recv_i, send_i, output_split_sizes=rs, input_split_sizes=split, group=group)
        pts, idx = recv_p, recv_i
        # the special last elements must sit last: the left maximum on the left range's last member, the pinned
        # element on the right range's last member (it arrives last anyway, its sender is the last source)
        if j == Gh - 1:
            k_ = composite(ordered_key(pts[:, axis]), idx)
            w = int((k_ == lmax[0]).nonzero()[0])
            last = pts.shape[0] - 1
            if w != last:
                sel = torch.arange(pts.shape[0], device=dev)
                sel[w], sel[last] = last, w
                pts, idx = pts[sel], idx[sel]
    # ---- local sub-tree by library sorts, then all_gather of the slices ----
    lo, n = seg_range(n_total, levels, rank)
    assert pts.shape[0] == n, (pts.shape[0], n)
    order = sort_build_local(pts, idx, levels)
    piece = idx[order]
    sizes = [seg_range(n_total, levels, r)[1] for r in range(world)]
    width = max(sizes)
    padded = torch.full((width,), -1, dtype=torch.int64, device=dev)
    padded[:n] = piece
    gathered = torch.empty((world * width,), dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    for r in range(world):
        rlo, rn = seg_range(n_total, levels, r)
        perm[rlo:rlo + rn] = gathered[r * width:r * width + rn]
    return perm.to(torch.int32)


def _lsr(x: torch.Tensor, s: int) -> torch.Tensor:
    """Logical shift right of int64 bit patterns."""
    if s == 0:
        return x
    if s >= 64:
        return torch.zeros_like(x)
    return (x >> s) & ((1 << (64 - s)) - 1)


def _group_all_reduce(t, gb, G, world, group):
    """Sum over the G consecutive ranks starting at gb, with world-wide collectives (one per step)."""
    if G == world:
        dist.all_reduce(t, group=group)
        return t
    buf = torch.zeros((world // G,) + tuple(t.shape), dtype=t.dtype, device=t.device)
    buf[gb // G] = t
    dist.all_reduce(buf, group=group)
    return buf[gb // G]


def _group_max(t, gb, G, world, group, full=False):
    lowest = torch.iinfo(t.dtype).min
    buf = torch.full((world // G,) + tuple(t.shape), lowest, dtype=t.dtype, device=t.device)
    buf[gb // G] = t
    dist.all_reduce(buf, op=dist.ReduceOp.MAX, group=group)
    return buf if full else buf[gb // G]


def _group_all_gather(t, gb, G, world, group):
    buf = torch.empty((world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(buf.view(-1), t.contiguous().view(-1), group=group)
    return buf[gb:gb + G]
