import sys, os, numpy as np
sys.path.insert(0, os.getcwd()); sys.path.insert(0, "tests")
from legged_mpc_control_b200 import abi, presets, synth
from legged_mpc_control_b200.engine import ConvexMpcEngine
from oracle import binding as ob
from parity_util import elem_err, norm_err, active_mismatch, row_residuals
p = presets.preset("northstar_a1")
B = 65536
b = synth.make_batch(B, p.horizon, p.dt, seed=4, gaits=(abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK), per_problem_params=True, base_params=p, gait_speed=(2.0, 4.5))
eng = ConvexMpcEngine(p, B)
got = eng.compute_grfs(b)
ref = ob.solve_batch(p, b, eliminate_swing=True, want_kkt=True)
print("status", np.bincount(got["status"], minlength=4))
for k in ("grf", "U"):
    ee = elem_err(got[k], ref[k]); ne = norm_err(got[k], ref[k])
    print(k, "elem q", np.quantile(ee, [0.5, 0.99, 0.999, 0.9999, 1.0]), "norm q", np.quantile(ne, [0.5, 0.99, 0.999, 0.9999, 1.0]))
print("active", active_mismatch(p, got["active_set"], ref["active_set"], ref["U"], b["mu"]))
ee = elem_err(got["U"], ref["U"])
np.set_printoptions(precision=6, linewidth=220)
for i in np.argsort(-ee)[:6]:
    pp = p.copy()
    for j in range(12): pp.q[j] = b["q"][i, j]; pp.r[j] = b["r"][i, j]
    pp.mu = b["mu"][i]
    one = {k: b[k][i:i+1] for k in ("x0", "rot", "foot_pos_abs", "cmd", "contact")}
    Hm, g, _ = ob.condensed_qp(pp, one, 0)
    c = b["contact"][i].reshape(-1).astype(bool)
    free = np.repeat(c, 3)
    Hr = Hm[np.ix_(free, free)]; gr = g[free]
    ug, ur = got["U"][i][free], ref["U"][i][free]
    f = lambda u: 0.5 * u @ Hr @ u + gr @ u
    ev = np.linalg.eigvalsh(Hr)
    rg = row_residuals(pp, got["U"][i:i+1])[0]; rr = row_residuals(pp, ref["U"][i:i+1])[0]
    print(f"--- {i} status {got['status'][i]} iters {got['iters'][i] & 0xffff},{got['iters'][i] >> 16} ee {ee[i]:.3e} ne {norm_err(got['U'], ref['U'])[i]:.3e} cond {ev[-1]/ev[0]:.3e} kkt {ref['kkt'][i]}")
    print("   obj got - obj ref", f(ug) - f(ur), "rel", (f(ug) - f(ur)) / abs(f(ur)), " min resid got", rg[:16*p.horizon].min(), rg[16*p.horizon:].min(), "ref", rr[:16*p.horizon].min(), rr[16*p.horizon:].min())
    d = np.abs(got["U"][i] - ref["U"][i]); j = d.argmax()
    print("   worst comp", j, got["U"][i][j], ref["U"][i][j], "active diff bits", np.nonzero(np.unpackbits((got["active_set"][i] ^ ref["active_set"][i]).view(np.uint8), bitorder="little"))[0])
