"""Execution runtimes: where shards live and how collectives between them run.

Two runtimes share every plan in this package:

* ``DistributedRuntime`` -- the production layout: ONE PROCESS PER GPU, rank r owns device r,
  collectives are ``torch.distributed`` (NCCL over NVLink 5 / NVSwitch; gloo in CPU tests).
  Replaces XLA:CPU's in-process collectives between the virtual devices that
  ``XLA_FLAGS=--xla_force_host_platform_device_count=8`` creates (case1a.py:3).
* ``VirtualRuntime`` -- N virtual devices inside one process whose buffers all live on one torch
  device (``cuda:0`` on a GPU box, ``cpu`` for plumbing tests).  This is the analogue of the
  reference's forced-host-device trick: the literal 8-device case scripts run on a single GPU,
  per-shard compute still goes through the C ABI, and "collectives" are plain buffer
  copies/adds between the virtual devices.

Per-shard compute is ALWAYS ``capi`` (libb200sl.so, CUDA only).  ``set_local_compute_for_testing``
exists so that CPU-only tests of the plans can inject a checker; nothing in the package does.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional, Sequence

import torch

from .sharding import Device

_runtime = None
_compute_override = None


def get_runtime():
    if _runtime is None:
        raise RuntimeError("no runtime: call init_virtual(n) or init_distributed() first")
    return _runtime


def set_runtime(rt):
    global _runtime
    _runtime = rt
    return rt


def init_virtual(num_devices: int, torch_device: Optional[str] = None):
    """N virtual devices in this process; buffers on ``torch_device`` (default: cuda:0 if present)."""
    if torch_device is None:
        torch_device = "cuda:0" if torch.cuda.is_available() else "cpu"
    return set_runtime(VirtualRuntime(num_devices, torch.device(torch_device)))


def init_distributed(backend: Optional[str] = None):
    """One process per GPU; reads RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* from the environment."""
    import torch.distributed as dist
    if not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        kwargs = {}
        if backend == "nccl":
            kwargs["device_id"] = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group(backend=backend, **kwargs)
    return set_runtime(DistributedRuntime())


def local_compute():
    """The per-shard compute provider: the CUDA C ABI, unless a test injected a checker."""
    if _compute_override is not None:
        return _compute_override
    from . import capi
    return capi


def set_local_compute_for_testing(provider):
    """TESTS ONLY: route per-shard compute of CPU-resident shards to ``provider`` (an object with the
    ``capi`` function names).  Pass None to restore the CUDA C ABI."""
    global _compute_override
    _compute_override = provider


class VirtualRuntime:
    kind = "virtual"

    def __init__(self, num_devices: int, torch_device: torch.device):
        self.torch_device = torch_device
        self.devices: List[Device] = [Device(i, "cuda" if torch_device.type == "cuda" else "host")
                                      for i in range(num_devices)]
        self.local_device_ids = list(range(num_devices))
        self.collective_log: List[tuple] = []     # (name, group size, bytes per member)

    def buffer_device(self, device_id: int) -> torch.device:
        return self.torch_device

    def _log(self, name, groups, bufs):
        g = groups[0]
        if len(g) > 1:   # one entry per collective call: (name, group size, bytes per member)
            self.collective_log.append((name, len(g), bufs[g[0]].numel() * bufs[g[0]].element_size()))

    # Every collective takes {device_id: tensor} for the addressable members and a list of groups
    # (each a list of device ids, ordered); returns {device_id: tensor}.
    def all_reduce(self, bufs: Dict[int, torch.Tensor], groups: Sequence[Sequence[int]]):
        out = {}
        for g in groups:
            total = bufs[g[0]].clone()
            for d in g[1:]:
                total += bufs[d]
            for d in g:
                out[d] = total if d == g[0] else total.clone()
        self._log("all_reduce", groups, bufs)
        return out

    def all_gather(self, bufs, groups, dim: int):
        out = {}
        for g in groups:
            cat = torch.cat([bufs[d] for d in g], dim=dim)
            for d in g:
                out[d] = cat if d == g[0] else cat.clone()
        self._log("all_gather", groups, bufs)
        return out

    def reduce_scatter(self, bufs, groups, dim: int):
        out = {}
        for g in groups:
            total = bufs[g[0]].clone()
            for d in g[1:]:
                total += bufs[d]
            for i, d in enumerate(g):
                out[d] = total.chunk(len(g), dim=dim)[i].contiguous()
        self._log("reduce_scatter", groups, bufs)
        return out

    def all_to_all(self, bufs, groups, split_dim: int, concat_dim: int):
        out = {}
        for g in groups:
            pieces = {d: bufs[d].chunk(len(g), dim=split_dim) for d in g}
            for i, d in enumerate(g):
                out[d] = torch.cat([pieces[s][i] for s in g], dim=concat_dim).contiguous()
        self._log("all_to_all", groups, bufs)
        return out

    def ring_shift_start(self, bufs, groups):
        """Every device's buffer moves to the NEXT member of its group (the last one's to the first): returns
        ({device: buffer received from the previous member}, pending).  In-process: nothing is pending."""
        out = {}
        for g in groups:
            for i, d in enumerate(g):
                out[d] = bufs[g[i - 1]]
        self._log("ring_shift", groups, bufs)
        return out, []

    def ring_shift_finish(self, pending):
        return None

    def exchange(self, bufs, src_index, wanted, global_shape, dtype):
        """Generic resharding: ``src_index[d]`` is the global slice device d holds (for all devices),
        ``wanted[d]`` the slice it must end up with.  In-process copy between virtual devices."""
        out = {}
        moved = 0
        for d, want in wanted.items():
            shape = tuple(s.stop - s.start for s in want)
            dst = torch.empty(shape, dtype=dtype, device=self.torch_device)
            covered = _fill_from_sources(dst, want, bufs, src_index)
            if not covered:
                raise RuntimeError(f"exchange: slice {want} for device {d} is not covered by the sources")
            moved += dst.numel() * dst.element_size()
            out[d] = dst
        self.collective_log.append(("exchange", len(wanted), moved // max(1, len(wanted))))
        return out

    def barrier(self):
        if self.torch_device.type == "cuda":
            torch.cuda.synchronize(self.torch_device)


def _intersect(a, b):
    out = []
    for x, y in zip(a, b):
        lo, hi = max(x.start, y.start), min(x.stop, y.stop)
        if lo >= hi:
            return None
        out.append(slice(lo, hi))
    return tuple(out)


def _fill_from_sources(dst, want, bufs, src_index) -> bool:
    """Copy into ``dst`` (global slice ``want``) every overlapping piece of the source shards.
    Returns True when the whole of ``want`` was covered."""
    # coverage is tracked on the destination's own device, whatever the size: a reshard whose sources cover the wanted
    # slice only partly must fail, never return uninitialised memory
    covered = torch.zeros(dst.shape, dtype=torch.bool, device=dst.device)
    for s, idx in src_index.items():
        if s not in bufs:
            continue
        inter = _intersect(want, idx)
        if inter is None:
            continue
        dsl = tuple(slice(i.start - w.start, i.stop - w.start) for i, w in zip(inter, want))
        ssl = tuple(slice(i.start - x.start, i.stop - x.start) for i, x in zip(inter, idx))
        dst[dsl].copy_(bufs[s][ssl])
        covered[dsl] = True
    return bool(covered.all())


class DistributedRuntime:
    kind = "distributed"

    def __init__(self):
        import torch.distributed as dist
        self.dist = dist
        self.rank = dist.get_rank()
        self.world = dist.get_world_size()
        backend = dist.get_backend()
        self.on_cuda = backend == "nccl"
        self.torch_device = torch.device("cuda", torch.cuda.current_device()) if self.on_cuda else torch.device("cpu")
        self.devices = [Device(i, "cuda" if self.on_cuda else "host") for i in range(self.world)]
        self.local_device_ids = [self.rank]
        self._groups = {}
        self.collective_log: List[tuple] = []

    def buffer_device(self, device_id: int) -> torch.device:
        return self.torch_device

    def group_for(self, ranks: Sequence[int]):
        """Process group of ``ranks``.  Groups are created collectively: every plan is computed
        identically on all ranks, so all ranks reach the same new_group calls in the same order."""
        key = tuple(ranks)
        if len(key) == self.world and list(key) == list(range(self.world)):
            return None   # default group
        if key not in self._groups:
            self._groups[key] = self.dist.new_group(ranks=list(key))
        return self._groups[key]

    def _my_group(self, groups):
        # create every group (collective call), return the one containing this rank
        mine = None
        for g in groups:
            pg = self.group_for(g) if len(g) > 1 else None
            if self.rank in g:
                mine = (list(g), pg)
        if mine is None:
            raise RuntimeError(f"rank {self.rank} is in none of the groups {groups}")
        return mine

    def all_reduce(self, bufs, groups):
        g, pg = self._my_group(groups)
        t = bufs[self.rank]
        if len(g) > 1:
            t = t.clone()
            self.dist.all_reduce(t, group=pg)
            self.collective_log.append(("all_reduce", len(g), t.numel() * t.element_size()))
        return {self.rank: t}

    def all_gather(self, bufs, groups, dim: int):
        g, pg = self._my_group(groups)
        t = bufs[self.rank].contiguous()
        if len(g) == 1:
            return {self.rank: t}
        parts = [torch.empty_like(t) for _ in g]
        self.dist.all_gather(parts, t, group=pg)
        self.collective_log.append(("all_gather", len(g), t.numel() * t.element_size()))
        return {self.rank: torch.cat(parts, dim=dim)}

    def reduce_scatter(self, bufs, groups, dim: int):
        g, pg = self._my_group(groups)
        t = bufs[self.rank]
        if len(g) == 1:
            return {self.rank: t}
        chunks = [c.contiguous() for c in t.chunk(len(g), dim=dim)]
        out = torch.empty_like(chunks[0])
        if self.on_cuda:
            self.dist.reduce_scatter(out, chunks, group=pg)
        else:   # gloo has no reduce_scatter: all_reduce then keep this rank's chunk
            full = t.clone()
            self.dist.all_reduce(full, group=pg)
            out = full.chunk(len(g), dim=dim)[g.index(self.rank)].contiguous()
        self.collective_log.append(("reduce_scatter", len(g), t.numel() * t.element_size()))
        return {self.rank: out}

    def all_to_all(self, bufs, groups, split_dim: int, concat_dim: int):
        g, pg = self._my_group(groups)
        t = bufs[self.rank]
        if len(g) == 1:
            return {self.rank: t}
        send = [c.contiguous() for c in t.chunk(len(g), dim=split_dim)]
        recv = [torch.empty_like(c) for c in send]
        if self.on_cuda:
            self.dist.all_to_all(recv, send, group=pg)
        else:   # gloo: emulate with all_gather of every member's pieces
            me = g.index(self.rank)
            for i, _ in enumerate(g):
                gathered = [torch.empty_like(send[i]) for _ in g]
                self.dist.all_gather(gathered, send[i], group=pg)
                if i == me:
                    recv = gathered
        self.collective_log.append(("all_to_all", len(g), t.numel() * t.element_size()))
        return {self.rank: torch.cat(recv, dim=concat_dim).contiguous()}

    def ring_shift_start(self, bufs, groups):
        """This rank's buffer goes to the next member of its group, the previous member's arrives: point-to-point
        send / receive, asynchronous -- the caller computes on what it holds and calls ring_shift_finish(pending) before
        touching the received buffer (ring attention: the transfer of block t+1 runs under the attention of block t)."""
        g = None
        for grp in groups:
            if self.rank in grp:
                g = list(grp)
        if g is None:
            raise RuntimeError(f"rank {self.rank} is in none of the groups {groups}")
        send = bufs[self.rank].contiguous()
        if len(g) == 1:
            return {self.rank: send}, []
        i = g.index(self.rank)
        recv = torch.empty_like(send)
        ops = [self.dist.P2POp(self.dist.isend, send, g[(i + 1) % len(g)]),
               self.dist.P2POp(self.dist.irecv, recv, g[(i - 1) % len(g)])]
        pending = self.dist.batch_isend_irecv(ops)
        self.collective_log.append(("ring_shift", len(g), send.numel() * send.element_size()))
        return {self.rank: recv}, (pending, send)          # `send` is kept alive until the transfer has completed

    def ring_shift_finish(self, pending):
        if pending:
            for req in pending[0]:
                req.wait()

    def exchange(self, bufs, src_index, wanted, global_shape, dtype):
        """Generic resharding over the whole job: all-gather every rank's shard (padded to a common
        size), then cut out the wanted slice locally.  Used only for small arrays / odd layouts;
        the large layers use the axis collectives above."""
        shapes = {d: tuple(s.stop - s.start for s in idx) for d, idx in src_index.items()}
        numel = max(int(torch.tensor(s).prod()) if len(s) else 1 for s in shapes.values())
        mine = bufs.get(self.rank)
        flat = torch.zeros(numel, dtype=dtype, device=self.torch_device)
        if mine is not None:
            flat[: mine.numel()] = mine.reshape(-1)
        parts = [torch.empty_like(flat) for _ in range(self.world)]
        self.dist.all_gather(parts, flat)
        self.collective_log.append(("exchange", self.world, numel * flat.element_size()))
        all_bufs = {}
        for d, shp in shapes.items():
            n = 1
            for s in shp:
                n *= s
            all_bufs[d] = parts[d][:n].reshape(shp)
        out = {}
        for d, want in wanted.items():
            if d != self.rank:
                continue
            shape = tuple(s.stop - s.start for s in want)
            dst = torch.empty(shape, dtype=dtype, device=self.torch_device)
            if not _fill_from_sources(dst, want, all_bufs, src_index):
                raise RuntimeError(f"exchange: slice {want} for device {d} is not covered by the sources")
            out[d] = dst
        return out

    def barrier(self):
        if self.on_cuda:
            torch.cuda.synchronize(self.torch_device)
        self.dist.barrier()
