"""How much of a step launch do the fused point operations cost?  C2 geometry, forward without store,
with the bench's 300 receivers / 2 sources against 1 receiver / 1 source (run under gpurun)."""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, psvwave_b200
for label, nr, both in (("300 receivers, 2 sources", 300, True), ("1 receiver, 2 sources", 1, True), ("1 receiver, 1 source", 1, False)):
    geom = dict(bench.GEOM, nr=nr, nt=400)
    cfg = bench.workload_config(400, 1, geom)
    if not both:
        cfg.update(n_sources=1, ix_sources=cfg["ix_sources"][:1], iz_sources=cfg["iz_sources"][:1],
                   moment_angles=cfg["moment_angles"][:1], which_source_to_fire_in_which_shot=[[0]])
    ini = os.path.join(tempfile.mkdtemp(), "p.ini")
    bench.write_ini(ini, cfg)
    m = psvwave_b200.fdModel(ini)
    m.set_parameter_fields(*bench.marmousi_like(m.nx, m.nz))
    m.forward_simulate(0, False)
    m.forward_simulate(0, False)
    print(label, "us/step %.2f" % (m._get_stat("elapsed_ms") * 1e3 / m.nt), flush=True)
