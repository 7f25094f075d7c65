"""Device meshes and sharding descriptions -- the host-side plumbing the reference scripts
reach through ``jax.sharding`` / ``jax.experimental.mesh_utils``:

  mesh_utils.create_device_mesh((2,4)) + PositionalSharding   case1a.py:15 (and case1b/2/3/4)
  PositionalSharding.reshape / .replicate(axis, keepdims)      case1a.py:24,30; case4_gspmd_ff.py:46,49
  Mesh(devices, ('data','model')), NamedSharding, PartitionSpec   case6_attention.py:155-161

A sharding answers one question for the rest of the library: "which global slice of an array
does device d hold?" (``devices_indices_map``).  All collective plans are derived from those
slices, never from axis names alone, so that permuted device orders such as case1a's
``sharding.reshape(4,2)`` work.
"""
from __future__ import annotations

import itertools
import threading
from dataclasses import dataclass
from typing import Dict, Optional, Sequence, Tuple

import numpy as np


@dataclass(frozen=True)
class Device:
    """One accelerator of the job.  ``id`` is the global index (= rank in a one-process-per-GPU
    job, = virtual device index in a single-process virtual mesh)."""
    id: int
    platform: str = "cuda"

    def __repr__(self):
        return f"{self.platform}:{self.id}"


def create_device_mesh(mesh_shape: Sequence[int], devices: Optional[Sequence[Device]] = None) -> np.ndarray:
    """Row-major ndarray of devices: device id = ravel index (id = 4*x + y on a (2,4) mesh).

    Mirrors ``mesh_utils.create_device_mesh`` for a single NVSwitch box, where every GPU is one
    hop from every other and no topology-aware permutation is needed (case1a.py:15)."""
    n = int(np.prod(mesh_shape))
    if devices is None:
        from . import runtime
        devices = runtime.get_runtime().devices
    if len(devices) < n:
        raise ValueError(f"mesh {tuple(mesh_shape)} needs {n} devices, job has {len(devices)}")
    arr = np.empty(n, dtype=object)
    for i in range(n):
        arr[i] = devices[i]
    return arr.reshape(tuple(mesh_shape))


class PartitionSpec(tuple):
    """Per-dimension mesh-axis names: ``PartitionSpec('data', 'model')``; ``None`` = replicated
    on that dim; a tuple of names splits one dim over several axes; may be shorter than the
    array rank (trailing dims replicated) -- case6_attention.py:161."""

    def __new__(cls, *partitions):
        return super().__new__(cls, partitions)

    def __repr__(self):
        return "PartitionSpec(" + ", ".join(repr(p) for p in self) + ")"


_mesh_stack = threading.local()


class Mesh:
    """Named view of a device ndarray; also a context manager like ``with mesh:``
    (case6_attention.py:155-156,234)."""

    def __init__(self, devices: np.ndarray, axis_names: Sequence[str]):
        devices = np.asarray(devices, dtype=object)
        if devices.ndim != len(axis_names):
            raise ValueError(f"mesh has {devices.ndim} dims but {len(axis_names)} axis names")
        self.devices = devices
        self.axis_names = tuple(axis_names)

    @property
    def shape(self) -> Dict[str, int]:
        return dict(zip(self.axis_names, self.devices.shape))

    @property
    def size(self) -> int:
        return int(self.devices.size)

    def axis_size(self, name: str) -> int:
        return self.devices.shape[self.axis_names.index(name)]

    def coords(self, device_id: int) -> Dict[str, int]:
        for idx in np.ndindex(self.devices.shape):
            if self.devices[idx].id == device_id:
                return dict(zip(self.axis_names, idx))
        raise KeyError(f"device {device_id} is not in the mesh")

    def axis_groups(self, axes: Sequence[str]) -> list:
        """Device-id groups that differ only in their coordinates on ``axes`` (the communicator
        of those mesh axes), members ordered by their coordinates on ``axes``."""
        axes = tuple(axes)
        other = [a for a in self.axis_names if a not in axes]
        groups = {}
        for idx in np.ndindex(self.devices.shape):
            c = dict(zip(self.axis_names, idx))
            key = tuple(c[a] for a in other)
            groups.setdefault(key, []).append((tuple(c[a] for a in axes), self.devices[idx].id))
        return [[d for _, d in sorted(v)] for _, v in sorted(groups.items())]

    def __enter__(self):
        stack = getattr(_mesh_stack, "stack", None)
        if stack is None:
            stack = _mesh_stack.stack = []
        stack.append(self)
        return self

    def __exit__(self, *exc):
        _mesh_stack.stack.pop()
        return False

    def __repr__(self):
        return f"Mesh({dict(self.shape)})"


def current_mesh() -> Optional[Mesh]:
    stack = getattr(_mesh_stack, "stack", None)
    return stack[-1] if stack else None


def _block(dim: int, nparts: int, part: int) -> slice:
    if dim % nparts != 0:
        raise ValueError(f"dimension of size {dim} is not divisible into {nparts} shards")
    step = dim // nparts
    return slice(part * step, (part + 1) * step)


class Sharding:
    """Base: maps every device of the sharding to the global index (tuple of slices) it holds."""

    def devices_indices_map(self, global_shape: Tuple[int, ...]) -> Dict[int, Tuple[slice, ...]]:
        raise NotImplementedError

    @property
    def device_ids(self) -> list:
        raise NotImplementedError

    def shard_shape(self, global_shape: Tuple[int, ...]) -> Tuple[int, ...]:
        idx = next(iter(self.devices_indices_map(tuple(global_shape)).values()))
        return tuple(s.stop - s.start for s in idx)

    def is_fully_replicated(self, global_shape) -> bool:
        return all(tuple(s.stop - s.start for s in idx) == tuple(global_shape)
                   for idx in self.devices_indices_map(tuple(global_shape)).values())

    def same_layout(self, other: "Sharding", global_shape) -> bool:
        a = self.devices_indices_map(tuple(global_shape))
        b = other.devices_indices_map(tuple(global_shape))
        return a.keys() == b.keys() and all(_slices_eq(a[d], b[d]) for d in a)


def _slices_eq(a, b) -> bool:
    return len(a) == len(b) and all(x.start == y.start and x.stop == y.stop for x, y in zip(a, b))


class NamedSharding(Sharding):
    """``NamedSharding(mesh, PartitionSpec(...))``; ``NamedSharding(mesh, None)`` = replicated
    (case6_attention.py:158-161,193)."""

    def __init__(self, mesh: Mesh, spec: Optional[PartitionSpec]):
        self.mesh = mesh
        self.spec = PartitionSpec() if spec is None else PartitionSpec(*spec)
        used = []
        for p in self.spec:
            for a in _axes_of(p):
                if a not in mesh.axis_names:
                    raise ValueError(f"unknown mesh axis {a!r} in {self.spec}")
                if a in used:
                    raise ValueError(f"mesh axis {a!r} used twice in {self.spec}")
                used.append(a)

    @property
    def device_ids(self):
        return [d.id for d in self.mesh.devices.flat]

    def dim_axes(self, ndim: int) -> list:
        """Tuple of mesh axes per array dim (padded with () up to ndim)."""
        out = [_axes_of(p) for p in self.spec]
        if len(out) > ndim:
            raise ValueError(f"{self.spec} has more entries than the array rank {ndim}")
        return out + [()] * (ndim - len(out))

    def devices_indices_map(self, global_shape):
        global_shape = tuple(global_shape)
        axes_per_dim = self.dim_axes(len(global_shape))
        out = {}
        for idx in np.ndindex(self.mesh.devices.shape):
            coord = dict(zip(self.mesh.axis_names, idx))
            sl = []
            for dim, axes in zip(global_shape, axes_per_dim):
                nparts, part = 1, 0
                for a in axes:   # major-to-minor over the listed axes
                    nparts *= self.mesh.axis_size(a)
                    part = part * self.mesh.axis_size(a) + coord[a]
                sl.append(_block(dim, nparts, part))
            out[self.mesh.devices[idx].id] = tuple(sl)
        return out

    def __repr__(self):
        return f"NamedSharding({self.mesh}, {self.spec})"


def _axes_of(p) -> tuple:
    if p is None:
        return ()
    if isinstance(p, str):
        return (p,)
    return tuple(p)


class PositionalSharding(Sharding):
    """``PositionalSharding(device_ndarray)``: the sharding's shape gives the number of pieces per
    array dim; ``.reshape`` re-views the same devices in row-major order; ``.replicate(axis,
    keepdims=True)`` collapses an axis to size 1 whose single piece is held by every device that
    was along it (case1a.py:15,24,30; case1b.py:24,30; case2.py:23,29; case3_fully_sharded.py:23,29;
    case4_gspmd_ff.py:42-49)."""

    def __init__(self, devices, _ids: Optional[np.ndarray] = None):
        if _ids is not None:
            self._ids = _ids                      # object ndarray: shape + (replicas,)
        else:
            devices = np.asarray(devices, dtype=object)
            self._ids = devices.reshape(devices.shape + (1,))

    @property
    def shape(self) -> Tuple[int, ...]:
        return self._ids.shape[:-1]

    @property
    def ndim(self) -> int:
        return self._ids.ndim - 1

    @property
    def device_ids(self):
        return [d.id for d in self._ids.flat]

    def reshape(self, *shape) -> "PositionalSharding":
        if len(shape) == 1 and isinstance(shape[0], (tuple, list)):
            shape = tuple(shape[0])
        return PositionalSharding(None, _ids=self._ids.reshape(tuple(shape) + (self._ids.shape[-1],)))

    def replicate(self, axis=None, keepdims: bool = True) -> "PositionalSharding":
        if axis is None:
            axes = tuple(range(self.ndim))
        elif isinstance(axis, int):
            axes = (axis % self.ndim,)
        else:
            axes = tuple(a % self.ndim for a in axis)
        ids = self._ids
        # move the replicated axes into the trailing replica dimension
        keep = [i for i in range(self.ndim) if i not in axes]
        perm = keep + list(axes) + [self.ndim]
        moved = np.transpose(ids, perm)
        new_shape = tuple(ids.shape[i] for i in keep) + (-1,)
        moved = moved.reshape(new_shape)
        if keepdims:
            full = [1 if i in axes else ids.shape[i] for i in range(self.ndim)] + [moved.shape[-1]]
            moved = moved.reshape(full)
        return PositionalSharding(None, _ids=moved)

    def devices_indices_map(self, global_shape):
        global_shape = tuple(global_shape)
        if len(global_shape) != self.ndim:
            raise ValueError(f"PositionalSharding of shape {self.shape} applied to an array of rank "
                             f"{len(global_shape)}")
        out = {}
        for idx in np.ndindex(self.shape):
            sl = tuple(_block(dim, n, i) for dim, n, i in zip(global_shape, self.shape, idx))
            for dev in self._ids[idx]:
                out[dev.id] = sl
        return dict(sorted(out.items()))

    def __repr__(self):
        return f"PositionalSharding(shape={self.shape}, replicas={self._ids.shape[-1]})"


class ReplicatedSharding(Sharding):
    """Every listed device holds the whole array."""

    def __init__(self, device_ids: Sequence[int]):
        self._device_ids = list(device_ids)

    @property
    def device_ids(self):
        return list(self._device_ids)

    def devices_indices_map(self, global_shape):
        full = tuple(slice(0, d) for d in global_shape)
        return {d: full for d in self._device_ids}


class SliceSharding(Sharding):
    """Explicit device -> global-slice table (what every plan ultimately works with)."""

    def __init__(self, table: Dict[int, Tuple[slice, ...]]):
        self._table = dict(sorted(table.items()))

    @property
    def device_ids(self):
        return list(self._table)

    def devices_indices_map(self, global_shape):
        return dict(self._table)
