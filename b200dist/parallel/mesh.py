"""N-D device mesh with named dimensions over process groups.

Same user-facing behaviour as ``torch.distributed.device_mesh.init_device_mesh`` as the reference
uses it (/root/reference/scripts/03_tensor_parallel_tp/01_device_mesh_basics.py:29,52-69;
scripts/06_hybrid_parallelism/01_fsdp_tp_hybrid.py:88-99): row-major rank layout, ``mesh["tp"]``
slicing, ``get_group`` / ``get_local_rank`` / ``size``.  Each 1-D slice also carries a
``b200dist.comm.Communicator`` (peer-memory data plane) created lazily.
"""
from __future__ import annotations

import itertools
from typing import Dict, List, Optional, Sequence, Tuple, Union

import torch
import torch.distributed as dist


class DeviceMesh:
    def __init__(self, device_type: str, shape: Sequence[int], mesh_dim_names: Optional[Sequence[str]] = None,
                 _ranks: Optional[torch.Tensor] = None, _parent: Optional["DeviceMesh"] = None,
                 _groups: Optional[Dict[int, object]] = None):
        self.device_type = device_type
        self.shape = tuple(int(s) for s in shape)
        self.mesh_dim_names = tuple(mesh_dim_names) if mesh_dim_names else tuple(f"dim{i}" for i in range(len(shape)))
        assert len(self.mesh_dim_names) == len(self.shape)
        world = dist.get_world_size() if dist.is_initialized() else 1
        n = 1
        for s in self.shape:
            n *= s
        if _ranks is None:
            assert n == world, f"mesh shape {self.shape} needs {n} ranks, world has {world}"
            _ranks = torch.arange(world).view(self.shape)
        self.mesh = _ranks
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self._parent = _parent
        self._groups: Dict[int, object] = _groups if _groups is not None else {}
        self._comms: Dict[int, object] = {}
        pos = (self.mesh == self.rank).nonzero()
        self._coord: Optional[Tuple[int, ...]] = tuple(int(x) for x in pos[0]) if pos.numel() else None
        if _groups is None:
            self._build_groups()

    # every rank creates every subgroup, in the same order (new_group is collective)
    def _build_groups(self) -> None:
        if not dist.is_initialized():
            return
        nd = len(self.shape)
        for d in range(nd):
            other = [range(s) for i, s in enumerate(self.shape) if i != d]
            for idx in itertools.product(*other):
                sl: List[Union[int, slice]] = list(idx)
                sl.insert(d, slice(None))
                ranks = [int(r) for r in self.mesh[tuple(sl)].tolist()]
                if nd == 1 and len(ranks) == dist.get_world_size() and ranks == list(range(len(ranks))):
                    g = dist.group.WORLD
                else:
                    g = dist.new_group(ranks)
                if self.rank in ranks:
                    self._groups[d] = g

    def _dim(self, dim: Union[int, str, None]) -> int:
        if dim is None:
            assert len(self.shape) == 1, "dim is required for an N-D mesh"
            return 0
        return self.mesh_dim_names.index(dim) if isinstance(dim, str) else int(dim)

    def __getitem__(self, name: Union[str, Tuple[str, ...]]) -> "DeviceMesh":
        """1-D sub-mesh along a named dim, containing this rank (``mesh["tp"]``)."""
        if isinstance(name, tuple):
            assert len(name) == 1, "only 1-D slices are supported"
            name = name[0]
        d = self._dim(name)
        sl: List[Union[int, slice]] = list(self._coord)
        sl[d] = slice(None)
        return DeviceMesh(self.device_type, (self.shape[d],), (self.mesh_dim_names[d],),
                          _ranks=self.mesh[tuple(sl)].clone(), _parent=self, _groups={0: self._groups.get(d)})

    def get_group(self, mesh_dim: Union[int, str, None] = None):
        return self._groups.get(self._dim(mesh_dim))

    def size(self, mesh_dim: Union[int, str, None] = None) -> int:
        if mesh_dim is None and len(self.shape) != 1:
            n = 1
            for s in self.shape:
                n *= s
            return n
        return self.shape[self._dim(mesh_dim)]

    def get_local_rank(self, mesh_dim: Union[int, str, None] = None) -> int:
        return self._coord[self._dim(mesh_dim)]

    def get_coordinate(self) -> Optional[Tuple[int, ...]]:
        return self._coord

    def ranks(self, mesh_dim: Union[int, str, None] = None) -> List[int]:
        d = self._dim(mesh_dim)
        sl: List[Union[int, slice]] = list(self._coord)
        sl[d] = slice(None)
        return [int(r) for r in self.mesh[tuple(sl)].tolist()]

    def communicator(self, mesh_dim: Union[int, str, None] = None):
        """Peer-memory communicator for this rank's slice along ``mesh_dim`` (collective on first use)."""
        from ..comm import get_communicator

        d = self._dim(mesh_dim)
        if d not in self._comms:
            g = self._groups.get(d)
            self._comms[d] = get_communicator(None if g is dist.group.WORLD else g)
        return self._comms[d]

    @property
    def ndim(self) -> int:
        return len(self.shape)

    def __repr__(self) -> str:
        return f"DeviceMesh({self.device_type}, shape={self.shape}, names={self.mesh_dim_names}, coord={self._coord})"


def init_device_mesh(device_type: str, mesh_shape: Sequence[int],
                     mesh_dim_names: Optional[Sequence[str]] = None) -> DeviceMesh:
    return DeviceMesh(device_type, mesh_shape, mesh_dim_names)
