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


class ExchangePlan:
    """Send / receive index lists of one rank.

    ``local_dst, local_src, local_code``  ghost cells whose source tile is owned by this rank;
    ``recv_dst[r], recv_code[r]``         ghost cells filled from peer r, in table order;
    ``send_src[r]``                       our interior cells peer r needs, in *its* table order.
    All offsets are element offsets into the local shard ``[tile_count, m, m]``.
    """

    def __init__(self, N, p, g, rank, ranges):
        self.N, self.p, self.g, self.rank, self.ranges = N, p, g, rank, list(ranges)
        n = N // p
        m = n + 2 * g
        tab = topology.build_halo_table(N, p, g)
        t0, tc = self.ranges[rank]
        owner_of = np.empty(topology.num_tiles(p), dtype=np.int64)
        for r, (b, c) in enumerate(self.ranges):
            owner_of[b:b + c] = r
        dst_owner = owner_of[tab["dst_tile"]]
        src_owner = owner_of[tab["src_tile"]]
        dst_off = (tab["dst_tile"] - t0) * m * m + tab["dst_j"] * m + tab["dst_i"]
        src_off_mine = (tab["src_tile"] - t0) * m * m + tab["src_j"] * m + tab["src_i"]
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
