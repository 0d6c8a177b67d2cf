"""Peer-mapped symmetric memory context for one process group (replaces the role of
``try_set_nki_parallel_state``, reference ``parallel_layers/parallel_state.py:425-444``).

Every rank allocates the same regions (``csrc/symm_mem.cu``: cudaMalloc + CUDA IPC), the handles are exchanged with
``all_gather_object`` over the group, and each rank maps all peers.  Kernels receive

  * ``data_ptrs[r]``   base of rank ``r``'s data region (own region at ``r == rank``),
  * ``sig_ptrs[r]``    base of rank ``r``'s signal pad (uint32 flags, zero-initialised),

as device arrays.  Flags are monotonically increasing epochs, never reset, so no clearing traffic is needed.
"""
from __future__ import annotations

import ctypes as _c
from typing import Dict, List, Tuple

import torch
import torch.distributed as dist

from . import _lib

SIGNAL_BYTES = 1 << 20


def _budget_bytes() -> int:
    """Upper bound on the symmetric memory of ONE context (``NXD_SYMM_BUDGET_MB``, default 16 GiB): regions are keyed
    by (kernel, shape) and live until the group is destroyed, so a program that keeps presenting new shapes would
    otherwise grow them without limit."""
    import os

    return int(os.environ.get("NXD_SYMM_BUDGET_MB", 16384)) << 20


class SymmMemContext:
    def __init__(self, group, ranks: List[int], data_bytes: int = 0):
        assert torch.cuda.is_available()
        self.group = group
        self.ranks = list(ranks)
        self.world = len(ranks)
        self.rank = dist.get_rank(group=group)
        self.device = torch.device("cuda", torch.cuda.current_device())
        self.lib = _lib.lib()
        self._regions: Dict[str, Tuple[int, List[int], int]] = {}
        self._epochs: Dict[str, int] = {}
        self._bytes: Dict[str, int] = {}          # region name -> bytes held on this rank (accounting, see memory_report)
        self.closed = False
        self.sig_local, self.sig_peers = self._alloc_exchange(SIGNAL_BYTES)
        self.sig_ptrs = torch.tensor(self.sig_peers, dtype=torch.int64, device=self.device)
        self._sig_cursor = 64   # uint32 slots [0, 64) are reserved for the barrier
        self.barrier()

    # -- allocation ------------------------------------------------------------------------------------------
    def _alloc_exchange(self, nbytes: int) -> Tuple[int, List[int]]:
        hb = self.lib.nxd_symm_handle_bytes()
        handle = (_c.c_ubyte * hb)()
        ptr = _c.c_void_p()
        _lib.check(self.lib.nxd_symm_alloc(_c.c_size_t(nbytes), _c.byref(ptr), handle), "nxd_symm_alloc")
        gathered = [None] * self.world
        dist.all_gather_object(gathered, bytes(handle), group=self.group)
        peers = []
        for r, h in enumerate(gathered):
            if r == self.rank:
                peers.append(ptr.value)
                continue
            buf = (_c.c_ubyte * hb).from_buffer_copy(h)
            p = _c.c_void_p()
            _lib.check(self.lib.nxd_symm_open(buf, _c.byref(p)), "nxd_symm_open")
            peers.append(p.value)
        return ptr.value, peers

    def region(self, name: str, nbytes: int) -> Tuple[int, torch.Tensor]:
        """Symmetric data region ``name`` of at least ``nbytes`` on every rank (collective on first use or growth).
        Returns ``(local_ptr, peer_ptr_table[int64 world])``."""
        cur = self._regions.get(name)
        if cur is None or cur[2] < nbytes:
            size = max(int(nbytes), 1 << 20)
            size = (size + (1 << 21) - 1) >> 21 << 21
            self._charge(name, size)
            local, peers = self._alloc_exchange(size)
            table = torch.tensor(peers, dtype=torch.int64, device=self.device)
            self._regions[name] = (local, peers, size, table)
            self.barrier()
        local, peers, size, table = self._regions[name]
        return local, table

    def region_mc(self, name: str, nbytes: int):
        """A symmetric region that also has an NVLS MULTICAST address (run at 2 and 8 GPUs), allocated and exchanged by ``torch.distributed._symmetric_memory`` (cuMem + cuMulticast under the hood; CUDA IPC memory from
        ``region`` cannot be bound to a multicast object).  Returns ``(local_ptr, peer_ptr_table, multicast_ptr)`` or
        ``None`` when the platform / group has no multicast support."""
        key = "mc:" + name
        cur = self._regions.get(key)
        if cur is not None and cur[2] >= nbytes:
            return cur[0], cur[3], cur[4]
        if getattr(self, "_mc_unavailable", False):     # asked once, answered once: do not retry on every call
            return None
        try:
            import torch.distributed._symmetric_memory as tsm

            size = (max(int(nbytes), 1 << 21) + (1 << 21) - 1) >> 21 << 21
            self._charge(key, size)
            buf = tsm.empty(size, dtype=torch.uint8, device=self.device)
            hdl = tsm.rendezvous(buf, self.group)
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            if not mc:
                self._mc_unavailable = True
                return None
            peers = [int(p) for p in hdl.buffer_ptrs]
            table = torch.tensor(peers, dtype=torch.int64, device=self.device)
            buf.zero_()
            # (local, peers, size, table, multicast, keep-alive references)
            self._regions[key] = (peers[self.rank], peers, size, table, mc, (buf, hdl))
            self.barrier()
            return peers[self.rank], table, mc
        except Exception as e:  # noqa: BLE001 - old driver, no NVSwitch, group not enabled for symmetric memory, ...
            from ..utils.logger import get_logger

            get_logger().warning("NVLS multicast region unavailable (%s): using per-peer pushes", e)
            self._mc_unavailable = True
            return None

    def _charge(self, name: str, size: int) -> None:
        """Account ``size`` bytes to region ``name`` (a grown region keeps its old allocation mapped by the peers until
        teardown, so growth adds up) and refuse to pass the budget with a message that names the consumers."""
        total = sum(self._bytes.values()) + size
        if total > _budget_bytes():
            raise RuntimeError(
                f"symmetric memory budget exceeded: {total >> 20} MiB > {_budget_bytes() >> 20} MiB "
                f"(NXD_SYMM_BUDGET_MB) while allocating {size >> 20} MiB for '{name}'.  " + self.memory_report())
        self._bytes[name] = self._bytes.get(name, 0) + size

    def memory_report(self) -> str:
        """One line per region, largest first: what holds the peer-mapped memory of this group on this rank."""
        rows = sorted(self._bytes.items(), key=lambda kv: -kv[1])
        return f"{sum(self._bytes.values()) >> 20} MiB in {len(rows)} regions: " + ", ".join(
            f"{n}={b >> 20}MiB" for n, b in rows[:12])

    def tensor_view(self, ptr: int, shape, dtype: torch.dtype) -> torch.Tensor:
        """Zero-copy torch tensor over symmetric memory at ``ptr`` (through ``__cuda_array_interface__``)."""
        nbytes = torch.empty(0, dtype=dtype).element_size()
        n = 1
        for d in shape:
            n *= int(d)

        class _Holder:
            __cuda_array_interface__ = {"shape": (n * nbytes,), "typestr": "|u1", "data": (int(ptr), False), "version": 2,
                                        "strides": None}

        raw = torch.as_tensor(_Holder(), device=self.device)
        return raw.view(dtype).view(tuple(int(d) for d in shape))

    def signal_slots(self, name: str, count: int) -> int:
        """Reserve ``count`` uint32 flags in the signal pad (same offset on every rank); returns the slot index."""
        key = "sig:" + name
        if key not in self._epochs:
            self._epochs[key] = self._sig_cursor
            self._sig_cursor += (count + 31) // 32 * 32
            assert self._sig_cursor * 4 <= SIGNAL_BYTES, "signal pad exhausted"
        return self._epochs[key]

    def next_epoch(self, name: str) -> int:
        e = self._epochs.get(name, 0) + 1
        self._epochs[name] = e
        return e

    def barrier(self) -> None:
        """Device-side rendezvous through the signal pads (all ranks of the group, current stream)."""
        e = self.next_epoch("__barrier__")
        _lib.check(self.lib.nxd_symm_barrier(_c.c_void_p(self.sig_ptrs.data_ptr()), _c.c_void_p(self.sig_local),
                                             _c.c_int(self.world), _c.c_int(self.rank), _c.c_int(0), _c.c_uint(e),
                                             _lib.stream_ptr()), "nxd_symm_barrier")

    def close(self) -> None:
        if self.closed:
            return
        self.closed = True
        torch.cuda.synchronize()
        for key, val in list(self._regions.items()):
            if key.startswith("mc:"):
                continue          # owned by torch.distributed._symmetric_memory (freed with the tensors we hold)
            local, peers = val[0], val[1]
            for r, p in enumerate(peers):
                if r != self.rank:
                    self.lib.nxd_symm_close(_c.c_void_p(p))
            self.lib.nxd_symm_free(_c.c_void_p(local))
        for r, p in enumerate(self.sig_peers):
            if r != self.rank:
                self.lib.nxd_symm_close(_c.c_void_p(p))
        self.lib.nxd_symm_free(_c.c_void_p(self.sig_local))
        self._regions.clear()
