"""Diagnostic: where do the CUDA path and the oracle first differ?  (run on the GPU box)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import waterlily_jl_b200 as wl
from oracle import wl_oracle as wo

m = 32; R = m / 8
body = dict(kind="sphere", center=(m / 2 - 1,) * 3, radius=R)
kw = dict(nu=2 * R / 3700)
sim = wl.Simulation((2 * m, m, m), (1.0, 0, 0), 2 * R, body=wl.AutoBody(**body), **kw)
ref = wo.OracleSim((2 * m, m, m), (1.0, 0, 0), 2 * R, body=wo.make_body(**body), **kw)
for nm in ("mu0", "mu1", "V", "sigma"):
    a, b = getattr(sim.flow, nm), ref.field(nm)
    print(nm, "max diff", np.abs(a - b).max(), "n diff", int((a != b).sum()))
inner = (slice(1, -1),) * 3
for step in range(6):
    sim.sim_step(remeasure=False); ref.sim_step(remeasure=False)
    u, p = sim.flow.u, sim.flow.p
    ru, rp = ref.field("u"), ref.field("p")
    dp = p - rp
    print(f"step {step}: n={sim.pois.n[-2:]} {ref.n[-2:]} dt={sim.flow.Δt[-1]:.7f}/{ref.dts[-1]:.7f} "
          f"max|du|={np.abs(u-ru).max():.2e} max|dp|={np.abs(dp).max():.2e} max|p|={np.abs(rp).max():.2e} "
          f"mean dp={dp[inner].mean():.2e} max|dp-mean|={np.abs(dp[inner]-dp[inner].mean()).max():.2e}")
    t0, t1 = sim.trace(0), ref.trace(0)
    print("   r1 gpu", ["%.6e" % v for v in t0["r1"]], "\n   r1 cpu", ["%.6e" % v for v in t1["r1"]])
