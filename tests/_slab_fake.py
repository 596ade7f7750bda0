"""CPU test double for one x-slab (tests/test_slab_gloo.py): the slab call surface of the CUDA
model (psv_create_slab_from_ini ... psv_halo_unpack, include/psvwave_b200.h) over a TOY update in
numpy -- one field, u_new[x] = mix of u[x], u[x+-1], u[x+-4], z-neighbours + source, which consumes 4 halo rows
per step exactly like the fused P-SV step -- so that the host logic of psvwave_b200.slab.SlabDomain
(partition, exchange order, halo_steps arithmetic, seismogram merge, gathers) runs without a GPU.
Not a solver: physics parity lives in tests/test_gpu_slab.py.
"""
import numpy as np

from oracle.oracle import config_from_ini


def partition(nx, rank, nranks, halo):
    base, rem = divmod(nx, nranks)
    own0 = rank * base + min(rank, rem)
    own1 = own0 + base + (1 if rank < rem else 0)
    lo = halo if rank > 0 else 0
    hi = halo if rank < nranks - 1 else 0
    return own0 - lo, own1 + hi, lo, hi


class ToySlab:
    def __init__(self, ini, rank, nranks, halo_rows):
        c = config_from_ini(ini)
        np_b = int(c["np_boundary"])
        self.nx_global = int(c["nx_inner"]) + 2 * np_b
        self.nz = int(c["nz_inner"]) + 2 * np_b
        self.nt, self.n_shots, self.nr = int(c["nt"]), int(c["n_shots"]), int(c["nr"])
        self.snapshot_interval = int(c["snapshot_interval"])
        self.rx = np.array(c["ix_receivers"], dtype=int) + np_b
        self.rz = np.array(c["iz_receivers"], dtype=int) + np_b
        self.sx = np.array(c["ix_sources"], dtype=int) + np_b
        self.sz = np.array(c["iz_sources"], dtype=int) + np_b
        self.x0, self.x1, self.lo, self.hi = partition(self.nx_global, rank, nranks, halo_rows)
        self.nx = self.x1 - self.x0
        self.H = max(self.lo, self.hi)
        self.u = np.zeros((self.nx, self.nz))
        self.scale = np.ones((self.nx, self.nz))
        self.ux = np.zeros((self.n_shots, self.nr, self.nt))
        self.uz = np.zeros_like(self.ux)
        self.send = [np.zeros(self.H * self.nz) if h else None for h in (self.lo, self.hi)]
        self.recv = [np.zeros(self.H * self.nz) if h else None for h in (self.lo, self.hi)]
        self.it = None

    # ---- surface used by SlabDomain --------------------------------------------------------
    def _set_option(self, name, value):
        pass

    def _slab_info(self):
        return dict(nx_global=self.nx_global, x_offset=self.x0, nx_local=self.nx, halo_lo=self.lo,
                    halo_hi=self.hi, pitch=self.nz, is_slab=True)

    def get_receivers(self, in_units, include_boundary):
        return self.rx.astype(float), self.rz.astype(float)

    def set_parameter_fields(self, vp, vs, rho):
        self.scale = np.array(vp) / 2000.0

    def get_parameter_fields(self):
        return self.scale * 2000.0, self.scale, self.scale

    def get_kernels(self):
        return self.u, 2 * self.u, 3 * self.u

    def get_synthetic_data(self):
        return self.ux.copy(), self.uz.copy()

    def set_synthetic_data(self, ux, uz):
        self.ux[...], self.uz[...] = ux, uz

    def synchronize(self):
        pass

    def _march_begin(self, i_shot, adjoint, store):
        self.u[...] = 0.0
        self.shot, self.adjoint = i_shot, adjoint
        self.it = self.nt - 1 if adjoint else 0

    def _march_steps(self, n):
        done = 0
        while done < n and 0 <= self.it < self.nt:
            it = self.it
            for r in range(self.nr):  # sample before the update, running sum in time
                x = self.rx[r] - self.x0
                if 0 <= x < self.nx and not self.adjoint:
                    v = self.u[x, self.rz[r]]
                    self.ux[self.shot, r, it] = v + (self.ux[self.shot, r, it - 1] if it else 0.0)
                    self.uz[self.shot, r, it] = -self.ux[self.shot, r, it]
            p = np.zeros((self.nx + 8, self.nz))
            p[4:-4] = self.u
            new = self.scale * (0.3 * p[4:-4] + 0.2 * (p[:-8] + p[8:]) + 0.1 * (p[3:-5] + p[5:-3]))
            new += 0.05 * (np.roll(self.u, 1, axis=1) + np.roll(self.u, -1, axis=1))
            for s in range(len(self.sx)):
                x = self.sx[s] - self.x0
                if 0 <= x < self.nx:
                    new[x, self.sz[s]] += np.sin(0.1 * (it + 1) * (s + 1 + self.shot))
            self.u = new
            self.it += -1 if self.adjoint else 1
            done += 1
        return done

    def _march_end(self):
        self.it = None

    def _halo_pack(self):
        if self.lo:
            self.send[0][...] = self.u[self.lo:self.lo + self.H].ravel()
        if self.hi:
            self.send[1][...] = self.u[self.nx - self.hi - self.H:self.nx - self.hi].ravel()

    def _halo_unpack(self):
        if self.lo:
            self.u[:self.H] = self.recv[0].reshape(self.H, self.nz)
        if self.hi:
            self.u[self.nx - self.H:] = self.recv[1].reshape(self.H, self.nz)

    def halo_views(self):
        return self.send[0], self.send[1], self.recv[0], self.recv[1]
