"""Shot sharding across the GPUs of one box: one process per GPU (torchrun), medium
replicated, shots dealt round-robin, ONE all-reduce of the three Lame-basis sensitivity
kernels per gradient, scalar all-reduce of the misfit.

Why this is the whole multi-GPU story for this path: forward/adjoint runs of different shots
are independent; the only cross-shot coupling in the reference is the additive accumulation
of lambda/mu/density kernels (src/fdModel.cpp:832-858) and the misfit sum (:1128-1145).
``torch.distributed`` (NCCL over NVLink on GPUs, gloo in the CPU tests) is plumbing only.

The kernels are reduced *before* ``map_kernels_to_velocity`` (which is linear in them), so
every rank ends with the same full gradient -- the state a single-GPU run of all shots has.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import numpy as np


class _CudaView:
    """Zero-copy view of a device allocation for torch (``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, count: int):
        self.__cuda_array_interface__ = {
            "shape": (count,), "typestr": "<f8", "data": (ptr, False), "version": 3, "strides": None}


class ShotParallel:
    """Drive one ``fdModel`` per rank over its share of the shots.

    ``model`` is any object with the ``fdModel`` method surface.  For the reduction it must
    offer either ``_device_kernel_buffer()`` (the CUDA model: pointer, element count) or
    ``lame_kernel_views()`` (host arrays; used by the gloo CPU tests).
    """

    def __init__(self, model, group=None, rank: Optional[int] = None, world_size: Optional[int] = None):
        self.model = model
        self.group = group
        if rank is None or world_size is None:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                rank, world_size = dist.get_rank(group), dist.get_world_size(group)
            else:
                rank, world_size = 0, 1
        self.rank, self.world_size = int(rank), int(world_size)

    # fixed map -> deterministic results for a given world size
    def my_shots(self) -> List[int]:
        return shots_of_rank(self.model.n_shots, self.rank, self.world_size)

    def forward(self, store_fields: bool = True) -> None:
        for s in self.my_shots():
            self.model.forward_simulate(s, store_fields)

    def misfit(self) -> float:
        """Sum of the L2 misfits of all shots (each rank contributes its own shots)."""
        local = sum(self.model.calculate_l2_misfit_shot(s) for s in self.my_shots())
        return self._allreduce_scalar(local)

    def gradient(self, release_snapshots: bool = False) -> float:
        """forward (stored) -> misfit -> adjoint sources -> adjoint -> all-reduce -> map.

        Returns the global misfit; afterwards ``get_kernels`` / ``get_gradient_vector`` are
        identical on every rank.  Equivalent to the reference's run_model
        (src/fdModel.cpp:1291-1308) with the shot loops sharded.
        """
        m = self.model
        m.reset_kernels()
        local = 0.0
        for s in self.my_shots():
            m.forward_simulate(s, True)
            local += m.calculate_l2_misfit_shot(s)
            m.calculate_l2_adjoint_sources()
            m.adjoint_simulate(s)
            if release_snapshots:
                m.release_snapshots(s)
        self.allreduce_kernels()
        m.map_kernels_to_velocity()
        return self._allreduce_scalar(local)

    # ---- collectives ------------------------------------------------------------------
    def allreduce_kernels(self) -> None:
        if self.world_size == 1:
            return
        import torch
        import torch.distributed as dist
        if hasattr(self.model, "_device_kernel_buffer"):
            ptr, count = self.model._device_kernel_buffer()
            dev = torch.device("cuda", int(self.model.device))
            t = torch.as_tensor(_CudaView(ptr, count), device=dev)
            torch.cuda.synchronize(dev)
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            torch.cuda.synchronize(dev)
        else:
            for view in self.model.lame_kernel_views():
                t = torch.from_numpy(view)  # shares memory with the model's array
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)

    def _allreduce_scalar(self, value: float) -> float:
        if self.world_size == 1:
            return float(value)
        import torch
        import torch.distributed as dist
        backend = dist.get_backend(self.group)
        dev = torch.device("cuda", int(self.model.device)) if backend == "nccl" else torch.device("cpu")
        t = torch.tensor([value], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return float(t.item())


def shots_of_rank(n_shots: int, rank: int, world_size: int) -> List[int]:
    """Round-robin shot -> rank map."""
    return [s for s in range(int(n_shots)) if s % world_size == rank]
