"""Domain split for grids too large for one GPU: x-slabs with halo exchange.

The reference has no counterpart (one ``fdModel`` holds the whole grid, src/fdModel.h:209-252);
this is the "large single grids" row of SURVEY.md 8(e).  The grid is cut along x -- rows are
contiguous in z, so a halo is a block of whole rows -- and every slab is an ordinary solver object
over its own rows plus ``halo_rows`` rows of each neighbour (``psv_create_slab_from_ini``).
One fused time step consumes 4 halo rows, so ``halo_rows = 4 * k`` lets k steps run between two
exchanges (default k = 4: 16 rows, ~3 % redundant work on 1000-row slabs, 4x fewer exchanges).

Three transports, same results:
  * ONE process drives all slabs, one per GPU of the box (``devices=[0, 1, ...]``) or several on one
    GPU: the whole march is native (``psv_march_slabs``) -- every slab pulls its neighbours' edge rows
    straight out of their state arrays over NVLink (peer-to-peer loads by one copy kernel), ordered by
    CUDA events between the slabs' streams; no staging, no NCCL, no PyTorch, no host synchronisation
    inside the march.  This is the default whenever the slabs live in one process;
  * one process per GPU (torchrun): each rank holds ONE slab; halos travel as NCCL point-to-point
    sends over NVLink (``torch.distributed.batch_isend_irecv`` on views of the staging buffers);
  * ``native=False``: slabs inside one process with staged device copies (the test doubles of the CPU
    tests use this path).

What is exchanged per side and exchange: 5 fields x halo_rows x pitch doubles, one contiguous
message.  Snapshots, sensitivity kernels and receiver samples are pointwise and stay local;
seismograms are merged (each receiver from the slab that owns its row) after a forward march,
kernels are gathered on demand.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np

from .parallel import _CudaView


class SlabDomain:
    """The ``fdModel`` call surface over a split grid.  Arrays in and out have the whole grid's shape."""

    def __init__(self, ini: str, n_slabs: Optional[int] = None, halo_steps: int = 4, group=None,
                 rank: Optional[int] = None, world_size: Optional[int] = None, options: Optional[dict] = None,
                 slab_factory=None, async_exchange: bool = True, devices: Optional[List[int]] = None,
                 native: bool = True):
        if rank is None or world_size is None:
            import sys
            dist = sys.modules.get("torch.distributed")  # an initialised process group implies the import
            if dist is not None and dist.is_available() and dist.is_initialized():
                rank, world_size = dist.get_rank(group), dist.get_world_size(group)
            else:
                rank, world_size = 0, 1  # all slabs in this process: PyTorch is not needed
        self.group, self.rank, self.world_size = group, int(rank), int(world_size)
        self.halo_rows = 4 * int(halo_steps)
        self.async_exchange = bool(async_exchange)
        self._views_cache = None
        self._native = False
        set_device = None
        if slab_factory is None:
            from . import fdModel, set_default_device as set_device
            slab_factory = fdModel._slab
            self._native = bool(native)
        if self.world_size > 1:
            if n_slabs not in (None, self.world_size):
                raise ValueError("one slab per rank: n_slabs must equal the world size")
            self.n_slabs, mine = self.world_size, [self.rank]
        else:
            self.n_slabs = int(n_slabs or 1)
            mine = list(range(self.n_slabs))
        self.slab_ranks = mine
        if devices and self.world_size > 1:
            raise ValueError("devices= places the slabs of ONE process; with one process per GPU each rank uses its own")
        self.devices = [int(d) for d in devices] if devices else None
        self.slabs = []
        for r in mine:
            if self.devices and set_device is not None:
                set_device(self.devices[r % len(self.devices)])  # slab r lives on that GPU
            self.slabs.append(slab_factory(ini, r, self.n_slabs, self.halo_rows))
        if self.devices and set_device is not None:
            set_device(self.devices[0])
        for m in self.slabs:
            for k, v in (options or {}).items():
                m._set_option(k, int(v))
        self.info = [m._slab_info() for m in self.slabs]
        m0, i0 = self.slabs[0], self.info[0]
        self.nx, self.nz, self.nt = int(i0["nx_global"]), int(m0.nz), int(m0.nt)
        self.n_shots, self.nr = int(m0.n_shots), int(m0.nr)
        self.snapshot_interval = int(m0.snapshot_interval)
        self.misfit = 0.0
        rx = m0.get_receivers(False, True)[0]
        self._rec_rows = np.asarray(rx, dtype=np.int64)

    # ---- geometry ----------------------------------------------------------------------------
    def owned_rows(self, k: int):
        """(first, last+1) global rows slab k of this process owns, and the same in its local rows."""
        i = self.info[k]
        lo, hi = int(i["halo_lo"]), int(i["nx_local"]) - int(i["halo_hi"])
        x0 = int(i["x_offset"])
        return (x0 + lo, x0 + hi), (lo, hi)

    def _rows(self, k: int):
        i = self.info[k]
        return int(i["x_offset"]), int(i["x_offset"]) + int(i["nx_local"])

    # ---- medium / data -----------------------------------------------------------------------
    def set_parameter_fields(self, vp, vs, rho) -> None:
        vp, vs, rho = (np.ascontiguousarray(a, dtype=np.float64) for a in (vp, vs, rho))
        if vp.shape != (self.nx, self.nz) or vs.shape != vp.shape or rho.shape != vp.shape:
            raise ValueError("parameter fields must have the whole grid's shape (nx, nz)")
        for k, m in enumerate(self.slabs):
            a, b = self._rows(k)
            m.set_parameter_fields(vp[a:b], vs[a:b], rho[a:b])

    def set_observed_data(self, ux, uz) -> None:
        for m in self.slabs:
            m.set_observed_data(ux, uz)

    def get_synthetic_data(self):
        return self.slabs[0].get_synthetic_data()

    def get_observed_data(self):
        return self.slabs[0].get_observed_data()

    def calculate_l2_misfit(self) -> float:
        for m in self.slabs:  # seismograms are replicated after the merge
            m.calculate_l2_misfit()
        self.misfit = float(self.slabs[0].misfit)
        return self.misfit

    def calculate_l2_adjoint_sources(self) -> None:
        for m in self.slabs:
            m.calculate_l2_adjoint_sources()

    def reset_kernels(self) -> None:
        for m in self.slabs:
            m.reset_kernels()

    def map_kernels_to_velocity(self) -> None:
        for m in self.slabs:  # pointwise (src/fdModel.cpp:1164-1183)
            m.map_kernels_to_velocity()

    def release_snapshots(self, i_shot: int = -1) -> None:
        for m in self.slabs:
            m.release_snapshots(i_shot)

    def synchronize(self) -> None:
        for m in self.slabs:
            m.synchronize()

    # ---- the time loops ----------------------------------------------------------------------
    def forward_simulate(self, i_shot: int, store_fields: bool = True, verbose: bool = False,
                         output_wavefields: bool = False, omp_threads_override: int = 0) -> None:
        if output_wavefields:
            raise ValueError("output_wavefields is not available on a split grid")
        self._march(i_shot, False, store_fields)
        self._merge_seismograms(i_shot)

    def adjoint_simulate(self, i_shot: int, verbose: bool = False, omp_threads_override: int = 0) -> None:
        self._march(i_shot, True, False)

    def _march(self, i_shot: int, adjoint: bool, store: bool) -> None:
        self.exchanges = (self.nt - 1) // max(self.halo_rows // 4, 1)
        if self._native and self.world_size == 1:
            from . import march_slabs
            march_slabs(self.slabs, i_shot, adjoint, store)  # peer-to-peer halos, all inside the library
            return
        begun = []
        try:
            for m in self.slabs:
                m._march_begin(i_shot, adjoint, store)
                begun.append(m)
            k = self.halo_rows // 4
            remaining = self.nt
            while remaining > 0:
                n = min(k, remaining)
                for m in self.slabs:
                    if m._march_steps(n) != n:
                        raise RuntimeError("slab marches out of step")
                remaining -= n
                if remaining > 0:
                    self.exchange_halos()
        finally:
            for m in begun:  # only what was begun (a slab whose begin failed holds no march)
                m._march_end()

    def _views(self, m):
        """(send_lo, send_hi, recv_lo, recv_hi) as torch tensors over the staging buffers."""
        import torch
        if hasattr(m, "halo_views"):  # host arrays (CPU test doubles)
            return [None if a is None else torch.from_numpy(a) for a in m.halo_views()]
        p = m._halo_buffers()
        dev = torch.device("cuda", int(m.device))
        return [None if not q else torch.as_tensor(_CudaView(q, p[4]), device=dev) for q in p[:4]]

    def exchange_halos(self) -> None:
        """Owned edge rows of every slab -> halo rows of its neighbours (all 5 fields)."""
        if self.n_slabs == 1:
            return
        import torch
        for m in self.slabs:
            m._halo_pack()
        if self.world_size > 1 and self._stream_ordered():
            # NCCL transport, no host synchronisation: the sends/receives are ordered after the pack and
            # before the unpack through the model's own stream (torch sees it as an external stream)
            import torch.distributed as dist
            m = self.slabs[0]
            if self._views_cache is None:
                dev = torch.device("cuda", int(m.device))
                self._views_cache = (self._views(m), torch.cuda.ExternalStream(m._stream(), device=dev))
            (s_lo, s_hi, r_lo, r_hi), stream = self._views_cache
            with torch.cuda.stream(stream):
                for req in dist.batch_isend_irecv(self._p2p_ops(s_lo, s_hi, r_lo, r_hi)):
                    req.wait()  # stream-side wait, the host runs on
            m._halo_unpack()
            return
        for m in self.slabs:
            m.synchronize()  # staging buffers complete before another stream / process reads them
        views = [self._views(m) for m in self.slabs]
        cuda = any(v is not None and v.is_cuda for v in views[0])
        if self.world_size > 1:
            import torch.distributed as dist
            for req in dist.batch_isend_irecv(self._p2p_ops(*views[0])):
                req.wait()
        else:
            for k in range(len(self.slabs) - 1):
                lower, upper = views[k], views[k + 1]
                upper[2].copy_(lower[1])  # my upper edge -> the next slab's lower halo
                lower[3].copy_(upper[0])  # its lower edge -> my upper halo
        if cuda:
            torch.cuda.current_stream().synchronize()
        for m in self.slabs:
            m._halo_unpack()

    def _stream_ordered(self) -> bool:
        import torch.distributed as dist
        return (self.async_exchange and hasattr(self.slabs[0], "_stream")
                and dist.get_backend(self.group) == "nccl")

    def _p2p_ops(self, s_lo, s_hi, r_lo, r_hi):
        import torch.distributed as dist
        ops = []
        if s_lo is not None:
            ops += [dist.P2POp(dist.isend, s_lo, self.rank - 1, self.group),
                    dist.P2POp(dist.irecv, r_lo, self.rank - 1, self.group)]
        if s_hi is not None:
            ops += [dist.P2POp(dist.isend, s_hi, self.rank + 1, self.group),
                    dist.P2POp(dist.irecv, r_hi, self.rank + 1, self.group)]
        return ops

    # ---- assembling whole-grid results --------------------------------------------------------
    def _sum_over_ranks(self, a: np.ndarray) -> np.ndarray:
        if self.world_size == 1:
            return a
        import torch
        import torch.distributed as dist
        if dist.get_backend(self.group) == "nccl":
            t = torch.from_numpy(a).to(torch.device("cuda", int(self.slabs[0].device)))
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            return t.cpu().numpy()
        t = torch.from_numpy(a)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return a

    def _merge_seismograms(self, i_shot: int) -> None:
        """Every receiver's trace from the slab that owns its row; replicated to all slabs."""
        if self.n_slabs == 1:
            return
        total = None
        own = []
        for k, m in enumerate(self.slabs):
            ux, uz = (np.array(a) for a in m.get_synthetic_data())
            own.append((ux, uz))
            (g0, g1), _ = self.owned_rows(k)
            mask = ((self._rec_rows >= g0) & (self._rec_rows < g1)).astype(np.float64)[:, None]
            part = np.stack([ux[i_shot] * mask, uz[i_shot] * mask])
            total = part if total is None else total + part
        total = self._sum_over_ranks(np.ascontiguousarray(total))
        for (ux, uz), m in zip(own, self.slabs):
            ux[i_shot], uz[i_shot] = total[0], total[1]
            m.set_synthetic_data(ux, uz)

    def gather_grid(self, getter) -> np.ndarray:
        """Whole-grid (nx, nz) array from per-slab local arrays: owned rows only."""
        out = np.zeros((self.nx, self.nz))
        for k, m in enumerate(self.slabs):
            (g0, g1), (l0, l1) = self.owned_rows(k)
            out[g0:g1] = np.asarray(getter(m)).reshape(-1, self.nz)[l0:l1]
        return self._sum_over_ranks(out)

    def get_kernels(self):
        return tuple(self.gather_grid(lambda m, i=i: m.get_kernels()[i]) for i in range(3))

    def get_parameter_fields(self):
        return tuple(self.gather_grid(lambda m, i=i: m.get_parameter_fields()[i]) for i in range(3))

    def get_gradient_vector(self) -> np.ndarray:
        """Block sums of the velocity-basis kernels over the free window, x fastest, [vp | vs | rho]
        (src/fdModel.cpp:1459-1515), assembled from the gathered kernels."""
        m = self.slabs[0]
        np_b = int(m.np_boundary)
        x0, x1 = np_b + int(m.nx_inner_boundary), np_b + int(m.nx_inner) - int(m.nx_inner_boundary)
        z0, z1 = np_b + int(m.nz_inner_boundary), np_b + int(m.nz_inner) - int(m.nz_inner_boundary)
        npx, npz = int(m.basis_gridpoints_x), int(m.basis_gridpoints_z)
        nbx, nbz = (x1 - x0) // npx, (z1 - z0) // npz
        parts = []
        for k in self.get_kernels():
            w = k[x0:x1, z0:z1].reshape(nbx, npx, nbz, npz)
            blocks = w.transpose(0, 2, 1, 3).reshape(nbx, nbz, npx * npz)
            acc = np.zeros((nbx, nbz))
            for j in range(npx * npz):  # the reference's running sum, in its (x, z) order inside a block
                acc = acc + blocks[:, :, j]
            parts.append(acc.T.ravel())
        return np.concatenate(parts)
