"""Worker of tests/test_parallel_gloo.py: one rank of a world_size-N gloo job on CPU.

Runs psvwave_b200.parallel.ShotParallel -- the host-side shot sharding + all-reduce logic that
drives one fdModel per GPU -- over the CPU oracle (test infrastructure standing in for the CUDA
model, which needs a device), and writes this rank's results to <out>/rank<r>.npz.

    python tests/_gloo_worker.py <ini> <out_dir> <obs.npz>      (RANK, WORLD_SIZE, MASTER_* from the env)
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch.distributed as dist  # noqa: E402

from oracle.oracle import OracleModel  # noqa: E402
from psvwave_b200 import parallel  # noqa: E402  (pure Python: shot map + collectives)


class HostModel(OracleModel):
    """fdModel surface + the two additive calls ShotParallel uses, on the CPU oracle."""

    def calculate_l2_misfit_shot(self, s):
        dt = self.dt
        total = 0.0
        for syn, obs in (("rtf_ux", "rtf_ux_true"), ("rtf_uz", "rtf_uz_true")):
            d = self.array(obs)[s] - self.array(syn)[s]
            total += float(np.sum(0.5 * dt * d * d))
        return total

    def lame_kernel_views(self):
        return [self.array(n).reshape(-1) for n in ("lambda_kernel", "mu_kernel", "density_l_kernel")]

    def release_snapshots(self, s):
        pass


def main():
    ini, out, obs_path = sys.argv[1:4]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    obs = np.load(obs_path)
    m = HostModel(ini)
    m.set_parameter_fields(obs["vp"], obs["vs"], obs["rho"])
    m.set_observed_data(obs["obs_ux"], obs["obs_uz"])
    sp = parallel.ShotParallel(m)
    misfit = sp.gradient(release_snapshots=True)
    ux, uz = m.get_synthetic_data()
    np.savez(os.path.join(out, f"rank{rank}.npz"), misfit=misfit, gradient=m.get_gradient_vector(),
             shots=np.array(sp.my_shots()), ux=ux, uz=uz, lambda_kernel=m.array("lambda_kernel"),
             vs_kernel=m.array("vs_kernel"))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
