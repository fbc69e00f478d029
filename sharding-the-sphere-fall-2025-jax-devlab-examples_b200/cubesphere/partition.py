"""Host-side exchange plan for a block partition of the tile axis.

The reference shards axis 0 (tiles) over a 1-D device mesh with
NamedSharding(mesh, P('tiles')) and lets GSPMD insert the cross-device copies
for every ``.at[face].set(...)`` (reference JAX-DevLab-Examples.py:77-79,
192-195).  Here the copies are explicit: each rank gathers the interior cells
its peers' ghosts need into one send buffer, one all-to-all moves the bytes
(NCCL over NVLink), and the receiver scatters them into its ghosts applying the
seam's signed permutation.

This module is the NumPy statement of that plan.  The CUDA library builds the
same plan in C++ (csrc/cs_topology.cu, cs_build_peer_plan); tests check that the
two agree and exercise this one over gloo with world_size 2 on the CPU.
"""
from __future__ import annotations

import numpy as np

from . import topology


def block_ranges(num_tiles: int, world: int):
    """[(tile_begin, tile_count)] per rank; requires num_tiles % world == 0 (reference PY:50)."""
    if world < 1 or world > num_tiles:
        raise ValueError(f"{world} devices requested for {num_tiles} tiles: every device needs at least one tile")
    if num_tiles % world:
        valid = [d for d in range(1, num_tiles + 1) if num_tiles % d == 0]
        raise ValueError(f"{num_tiles} tiles cannot be split evenly over {world} devices; use one of {valid}")
    per = num_tiles // world
    return [(r * per, per) for r in range(world)]


def tile_adjacency(p: int):
    """[(tile_a, tile_b)] for the 12 p^2 tile edges of the sphere (each once, a < b)."""
    desc = topology.edge_push_descriptors(p)
    edges = set()
    for t in range(desc.shape[0]):
        for e in range(4):
            a, b = t, int(desc[t, e, 0])
            edges.add((min(a, b), max(a, b)))
    return sorted(edges)


def cut_edges(order, p: int, world: int):
    """Number of tile edges whose two tiles lie on different devices when axis 0, in this order, is block-partitioned."""
    n = topology.num_tiles(p)
    per = n // world
    owner = np.empty(n, dtype=np.int64)
    owner[np.asarray(order)] = np.arange(n) // per
    return sum(1 for a, b in tile_adjacency(p) if owner[a] != owner[b])


def locality_order(p: int, world: int, restarts: int = 64, seed: int = 20251118):
    """A permutation of the tiles along axis 0 whose plain block partition over ``world`` devices cuts few tile edges:
    the reference's ``P('tiles')`` contract (JAX-DevLab-Examples.py:77-79) is kept -- equal contiguous blocks of axis 0
    -- and only the ORDER of the tiles along the axis changes (the deck's "Graph Partitioner" box, slide 7).

    Deterministic local search (pairwise swaps between devices, random restarts with a fixed seed, ties broken by the
    smaller permutation), so every rank computes the same answer.  For 24 tiles on 8 devices it finds the obvious optimum:
    every device owns the three tiles around one cube vertex (24 of 48 edges cut instead of 36 for the block order, and
    three neighbour devices each instead of up to seven); on 4 devices 16 instead of 26.
    """
    n = topology.num_tiles(p)
    if n % world:
        raise ValueError(f"{n} tiles cannot be split evenly over {world} devices")
    per = n // world
    if world == 1 or per == 1:
        return np.arange(n, dtype=np.int32)
    edges = tile_adjacency(p)
    nbr = [[] for _ in range(n)]
    for a, b in edges:
        nbr[a].append(b)
        nbr[b].append(a)
    rng = np.random.default_rng(seed)

    def cut(owner):
        return sum(1 for a, b in edges if owner[a] != owner[b])

    def gain(owner, t, new):                    # change of the cut when tile t moves to device `new`
        old = owner[t]
        return sum(int(owner[q] == old) - int(owner[q] == new) for q in nbr[t])

    best_owner, best_cut = None, None
    for r in range(restarts):
        owner = [int(x) for x in np.arange(n) // per]
        if r:
            owner = [owner[k] for k in rng.permutation(n)]
        improved = True
        while improved:
            improved = False
            for a in range(n):
                for b in range(a + 1, n):
                    if owner[a] == owner[b]:
                        continue
                    oa, ob = owner[a], owner[b]
                    d = gain(owner, a, ob)
                    owner[a] = ob
                    d += gain(owner, b, oa)
                    owner[b] = oa
                    if d < 0:
                        improved = True
                    else:
                        owner[a], owner[b] = oa, ob
        c = cut(owner)
        # canonical labelling: devices numbered by their smallest tile, so equal partitions compare equal
        first = {}
        for t in range(n):
            first.setdefault(owner[t], len(first))
        canon = np.array([first[o] for o in owner])
        key = (c, tuple(canon))
        if best_cut is None or key < best_cut:
            best_cut, best_owner = key, canon
    order = np.concatenate([np.nonzero(best_owner == d)[0] for d in range(world)]).astype(np.int32)
    return order


class ExchangePlan:
    """Send / receive index lists of one rank.

    ``local_dst, local_src, local_code``  ghost cells whose source tile is owned by this rank;
    ``recv_dst[r], recv_code[r]``         ghost cells filled from peer r, in table order;
    ``send_src[r]``                       our interior cells peer r needs, in *its* table order.
    All offsets are element offsets into the local shard ``[tile_count, m, m]``.
    """

    def __init__(self, N, p, g, rank, ranges, order=None):
        """``ranges``: (begin, count) of every rank in POSITIONS along axis 0; ``order``: position -> tile id
        (default: the identity, i.e. the reference's block order)."""
        self.N, self.p, self.g, self.rank, self.ranges = N, p, g, rank, list(ranges)
        n = N // p
        m = n + 2 * g
        n_all = topology.num_tiles(p)
        order = np.arange(n_all) if order is None else np.asarray(order, dtype=np.int64)
        pos_of = np.empty(n_all, dtype=np.int64)
        pos_of[order] = np.arange(n_all)
        # ghost cells in the order the library enumerates them: position by position
        tab0 = topology.build_halo_table(N, p, g)
        per_tile = tab0["dst_tile"].size // n_all
        sel = np.concatenate([np.arange(t * per_tile, (t + 1) * per_tile) for t in order])
        tab = {k: v[sel] for k, v in tab0.items()}
        t0, tc = self.ranges[rank]
        owner_of = np.empty(n_all, dtype=np.int64)
        for r, (b, c) in enumerate(self.ranges):
            owner_of[order[b:b + c]] = r
        dst_owner = owner_of[tab["dst_tile"]]
        src_owner = owner_of[tab["src_tile"]]
        dst_off = (pos_of[tab["dst_tile"]] - t0) * m * m + tab["dst_j"] * m + tab["dst_i"]
        src_off_mine = (pos_of[tab["src_tile"]] - t0) * m * m + tab["src_j"] * m + tab["src_i"]
        mine = dst_owner == rank
        loc = mine & (src_owner == rank)
        self.local_dst, self.local_src, self.local_code = dst_off[loc], src_off_mine[loc], tab["code"][loc]
        self.recv_dst, self.recv_code, self.send_src = [], [], []
        for r in range(len(self.ranges)):
            sel = mine & (src_owner == r) & (r != rank)
            self.recv_dst.append(dst_off[sel])
            self.recv_code.append(tab["code"][sel])
            sel = (dst_owner == r) & (src_owner == rank) & (r != rank)
            self.send_src.append(src_off_mine[sel])
        self.send_counts = [int(x.size) for x in self.send_src]
        self.recv_counts = [int(x.size) for x in self.recv_dst]

    # NumPy emulation of the pack / unpack kernels (layout: per peer [field][cell])
    def pack(self, fields):
        return [np.stack([f.reshape(-1)[idx] for f in fields]) for idx in self.send_src]

    def unpack(self, blocks, fields):
        """``blocks[r]`` has shape [nfields, recv_counts[r]]; fields = (h,) or (h, u, v), updated in place."""
        for r, blk in enumerate(blocks):
            if self.recv_counts[r] == 0:
                continue
            dst, code = self.recv_dst[r], self.recv_code[r]
            fields[0].reshape(-1)[dst] = blk[0]
            if len(fields) == 3:
                a, b = apply_code(blk[1], blk[2], code)
                fields[1].reshape(-1)[dst] = a
                fields[2].reshape(-1)[dst] = b

    def fill_local(self, fields):
        fields[0].reshape(-1)[self.local_dst] = fields[0].reshape(-1)[self.local_src]
        if len(fields) == 3:
            a, b = apply_code(fields[1].reshape(-1)[self.local_src], fields[2].reshape(-1)[self.local_src],
                              self.local_code)
            fields[1].reshape(-1)[self.local_dst] = a
            fields[2].reshape(-1)[self.local_dst] = b


def apply_code(u, v, code):
    """Signed permutation of vector components (bit0 swap, bit1 negate out-u, bit2 negate out-v)."""
    swap = (code & 1) == 1
    a = np.where(swap, v, u)
    b = np.where(swap, u, v)
    return np.where(code & 2, -a, a), np.where(code & 4, -b, b)


class InterpExchangePlan:
    """Exchange plan of one rank for the SECOND-ORDER ghosts (geometry.seam_interp_tables) on a sharded sphere.

    Every ghost cell is a weighted sum of four interior cells of a neighbouring face; on a shard those four may live on
    this rank or on peers (and, with several tiles per face edge, on two different peers).  So the raw source cells
    travel and the receiver interpolates: for every peer r, ``need[r]`` = the peer's cells this rank's ghost rows read
    (sorted by global element offset, each once) and ``send_src[r]`` = this rank's cells peer r reads, in the order r
    expects them.  Rows address their sources in one virtual space: offsets below ``field_elems`` are cells of the local
    shard, ``field_elems + k`` is cell k of the receive buffer (peer blocks in rank order, 1 or 3 values per cell).
    Everything is derived from the global table, so all ranks compute consistent plans without talking.

    The NumPy methods emulate the device path (tests run all ranks in one process)."""

    def __init__(self, N, p, g, rank, ranges, order=None, points=4):
        from . import geometry
        self.N, self.p, self.g, self.rank, self.ranges = N, p, g, rank, list(ranges)
        n = N // p
        m = n + 2 * g
        te = m * m
        n_all = topology.num_tiles(p)
        order = np.arange(n_all) if order is None else np.asarray(order, dtype=np.int64)
        pos_of = np.empty(n_all, dtype=np.int64)
        pos_of[order] = np.arange(n_all)
        world = len(self.ranges)
        owner_of = np.empty(n_all, dtype=np.int64)
        for r, (b, c) in enumerate(self.ranges):
            owner_of[order[b:b + c]] = r
        tab = geometry.seam_interp_tables(N, g, p, points)
        self.tangents = tab["tangents"]
        dst_tile = tab["dst"] // te
        src_tile = tab["src"] // te                                  # [rows, 4]
        dst_owner = owner_of[dst_tile]
        src_owner = owner_of[src_tile]
        t0 = [b for b, _ in self.ranges]

        def local(glob, r):                                          # global element offset -> offset in rank r's shard
            return (pos_of[glob // te] - t0[r]) * te + glob % te

        # need[q][r]: global offsets of rank r's cells that rank q's rows read (sorted, unique)
        need = [[None] * world for _ in range(world)]
        for q in range(world):
            rows = dst_owner == q
            s, o = tab["src"][rows], src_owner[rows]
            for r in range(world):
                need[q][r] = np.unique(s[o == r]) if r != q else np.empty(0, dtype=np.int64)
        self.recv_counts = [int(need[rank][r].size) for r in range(world)]
        self.send_counts = [int(need[r][rank].size) for r in range(world)]
        self.send_src = [local(need[r][rank], rank) for r in range(world)]
        self.remote_offsets = [int(sum(need[r][q].size for q in range(rank))) for r in range(world)]
        self.field_elems = self.ranges[rank][1] * te
        mine = np.nonzero(dst_owner == rank)[0]
        self.dst = local(tab["dst"][mine], rank)
        src = tab["src"][mine]
        own = src_owner[mine]
        virt = np.empty_like(src)
        loc = own == rank
        virt[loc] = local(src[loc], rank)
        roff = np.concatenate([[0], np.cumsum(self.recv_counts)])
        for r in range(world):
            sel = own == r
            if r == rank or not sel.any():
                continue
            virt[sel] = self.field_elems + roff[r] + np.searchsorted(need[rank][r], src[sel])
        self.src = virt
        self.w, self.rot = tab["w"][mine], tab["rot"][mine]

    # NumPy emulation: cells [cell, field] per peer
    def pack(self, fields):
        return [np.stack([f.reshape(-1)[idx] for f in fields], axis=-1) for idx in self.send_src]

    def apply(self, blocks, fields):
        """``blocks[r]``: what peer r packed for this rank, ``[recv_counts[r], nfields]``; fills every ghost cell."""
        nf = len(fields)
        recv = np.concatenate([np.asarray(b).reshape(-1, nf) for b in blocks]) if sum(self.recv_counts) else np.zeros((0, nf))
        out = []
        for k, f in enumerate(fields):
            flat = f.reshape(-1)
            virt = np.concatenate([flat, recv[:, k]])
            out.append(np.sum(self.w * virt[self.src], -1))
        fields[0].reshape(-1)[self.dst] = out[0]
        if nf == 3:
            fields[1].reshape(-1)[self.dst] = self.rot[:, 0, 0] * out[1] + self.rot[:, 0, 1] * out[2]
            fields[2].reshape(-1)[self.dst] = self.rot[:, 1, 0] * out[1] + self.rot[:, 1, 1] * out[2]
