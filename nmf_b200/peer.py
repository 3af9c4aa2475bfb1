"""Peer-memory arena of a sharded Engine: one buffer per rank with the same byte layout everywhere, mapped into
every rank's address space, so that the ranks' own kernels store their factor windows, Gram terms and scalars
straight into the peers' memory over NVLink (csrc/peer.cu).

torch.distributed._symmetric_memory provides the allocation and the exchange of the handles (plumbing); the data
path is nmfb_peer_push / nmfb_peer_barrier / nmfb_sum_rank_slots / nmfb_adopt_exchange of the C ABI.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Tuple

import torch
import torch.distributed as dist

from ._lib import check

FLAG_BYTES = 256          # [world] uint32 barrier flags at offset 0 of every arena (world <= 16)


def _open_symmetric(nbytes: int, device, group) -> Tuple[torch.Tensor, List[int], object]:
    import torch.distributed._symmetric_memory as symm
    pg = group if group is not None else dist.group.WORLD
    try:                                     # older releases want the group enabled first; newer ones deprecate it
        symm.enable_symm_mem_for_group(pg.group_name)
    except Exception:                        # noqa: BLE001
        pass
    t = symm.empty(nbytes, dtype=torch.uint8, device=device)
    t.zero_()
    torch.cuda.current_stream(device).synchronize()
    hdl = symm.rendezvous(t, pg)             # collective: every rank has zeroed its arena before any pointer exists
    return t, [int(p) for p in hdl.buffer_ptrs], hdl


class PeerArena:
    """`layout`: list of (name, nbytes); every region is 256-byte aligned and starts at the same offset on all
    ranks.  Regions that are written by peers are double-buffered by their users (Engine)."""

    def __init__(self, lib, layout, device, rank: int, world: int, group=None):
        self.lib, self.device, self.rank, self.world, self.group = lib, device, rank, world, group
        self.offsets = {}
        off = FLAG_BYTES
        for name, nbytes in layout:
            self.offsets[name] = off
            off += (int(nbytes) + 255) // 256 * 256
        self.nbytes = off
        self.local, self.ptrs, self._handle = _open_symmetric(self.nbytes, device, group)
        assert len(self.ptrs) == world and self.ptrs[rank] == self.local.data_ptr()
        self._ptr_arr = (C.c_void_p * world)(*self.ptrs)
        self.seq = 0
        # NVSwitch multicast mapping of the arenas (one store, replicated by the switch into every rank), when the
        # symmetric-memory handle has one; NMFB_NO_MULTICAST=1 keeps the unicast stores (A/B measurements)
        mc = 0
        try:
            if bool(getattr(self._handle, "has_multicast_support", False)):
                mc = int(self._handle.multicast_ptr or 0)
        except Exception:                        # noqa: BLE001
            mc = 0
        self.mc_ptr = 0 if os.environ.get("NMFB_NO_MULTICAST", "0") == "1" else mc

    def view(self, name: str, shape, dtype) -> torch.Tensor:
        """Tensor view of the LOCAL copy of a region."""
        n = int(torch.tensor([], dtype=dtype).element_size())
        count = 1
        for s in shape:
            count *= int(s)
        off = self.offsets[name]
        return self.local[off: off + count * n].view(dtype).view(*shape)

    def push(self, src_ptr: int, src_pitch: int, region: str, dst_off: int, dst_pitch: int, rows: int,
             row_bytes: int, stream: int):
        """rows x row_bytes from src (device pointer, pitch in bytes) into EVERY rank's `region` at dst_off."""
        if self.mc_ptr:
            check(self.lib.nmfb_peer_push_mc(src_ptr, src_pitch, self.mc_ptr, self.offsets[region] + dst_off, dst_pitch,
                                             rows, row_bytes, stream))
            return
        check(self.lib.nmfb_peer_push(src_ptr, src_pitch, self._ptr_arr, self.world, self.offsets[region] + dst_off,
                                      dst_pitch, rows, row_bytes, stream))

    def scatter(self, src_ptr: int, src_pitch: int, src_peer_stride: int, region: str, dst_off: int, dst_pitch: int,
                rows: int, row_bytes, stream: int):
        """All-to-all: rank q's `region` receives rows x row_bytes[q] from src + q * src_peer_stride at dst_off."""
        arr = (C.c_long * self.world)(*[int(b) for b in row_bytes])
        check(self.lib.nmfb_peer_scatter(src_ptr, src_pitch, src_peer_stride, self._ptr_arr, self.world,
                                         self.offsets[region] + dst_off, dst_pitch, rows, arr, stream))

    def barrier(self, stream: int):
        """All ranks call this the same number of times, in the same order."""
        self.seq += 1
        check(self.lib.nmfb_peer_barrier(self._ptr_arr, self.rank, self.world, self.seq & 0xFFFFFFFF, stream))
