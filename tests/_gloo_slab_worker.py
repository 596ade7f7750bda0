"""Worker of tests/test_slab_gloo.py: one rank (= one slab) of a world_size-N gloo job on CPU.
Runs psvwave_b200.slab.SlabDomain over the numpy test double of tests/_slab_fake.py and writes the
whole-grid results this rank sees to <out>/rank<r>.npz.

    python tests/_gloo_slab_worker.py <ini> <out_dir> <halo_steps>   (RANK, WORLD_SIZE, MASTER_* from the env)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch.distributed as dist  # noqa: E402

from _slab_fake import ToySlab  # noqa: E402
from psvwave_b200.slab import SlabDomain  # noqa: E402


def run(domain):
    vp = 2000.0 * (1 + 0.01 * np.cos(np.arange(domain.nx * domain.nz).reshape(domain.nx, domain.nz)))
    domain.set_parameter_fields(vp, vp, vp)
    for s in range(domain.n_shots):
        domain.forward_simulate(s, True)
    ux, uz = domain.get_synthetic_data()
    field = domain.get_kernels()[1]
    domain.adjoint_simulate(0)
    return dict(ux=ux, uz=uz, field=field, adj=domain.get_kernels()[0], vp=domain.get_parameter_fields()[0])


def main():
    ini, out, halo_steps = sys.argv[1], sys.argv[2], int(sys.argv[3])
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    d = SlabDomain(ini, halo_steps=halo_steps, slab_factory=ToySlab)
    assert d.n_slabs == world and d.slab_ranks == [rank]
    np.savez(os.path.join(out, f"rank{rank}.npz"), owned=np.array(d.owned_rows(0)[0]), **run(d))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
