"""One or a few steps of a bench workload, for ncu captures: python scripts/r2/one_step.py [workload] [steps]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import bench
import __graft_entry__ as ge
ge.build()
import waterlily_jl_b200 as wl
name = sys.argv[1] if len(sys.argv) > 1 else "sphere"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
desc, dims, uBC, L, kw, body, remeasure = bench.workload(name)
sim = wl.Simulation(dims, uBC, L, body=wl.AutoBody(**body) if body else None, **kw)
for _ in range(steps):
    sim.sim_step(remeasure=remeasure)
sim.sync()
print(desc, sim.pois.n, sim.launch_count())
