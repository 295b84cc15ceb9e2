"""Small fixed workload for profiling (ncu): one device-resident solve of `B` problems.
Usage: python tools_profile.py [B] [preset]"""
import sys

import torch

from legged_mpc_control_b200 import presets, synth
from legged_mpc_control_b200.engine import ConvexMpcEngine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
name = sys.argv[2] if len(sys.argv) > 2 else "northstar_a1"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
p = presets.preset(name)
from legged_mpc_control_b200 import abi
go1 = "go1" in name
gaits = (abi.GAIT_TROT, abi.GAIT_BOUND, abi.GAIT_PRONK) if go1 else (abi.GAIT_TROT,)
if len(sys.argv) > 4:
    gaits = tuple({"trot": abi.GAIT_TROT, "stand": abi.GAIT_STAND, "crawl": abi.GAIT_CRAWL, "tws": abi.GAIT_TROT_WITH_STAND,
                   "bound": abi.GAIT_BOUND, "pronk": abi.GAIT_PRONK}[g] for g in sys.argv[4].split(","))
b = synth.make_batch(B, p.horizon, p.dt, seed=2 if go1 else 3, robot="go1" if go1 else "a1", gaits=gaits)
dev = torch.device("cuda:0")
tin = {k: torch.from_numpy(b[k]).to(dev) for k in ("x0", "rot", "foot_pos_abs", "cmd", "contact")}
grf = torch.zeros(B, 12, dtype=torch.float64, device=dev)
st = torch.zeros(B, dtype=torch.int32, device=dev)
it = torch.zeros(B, dtype=torch.int32, device=dev)
eng = ConvexMpcEngine(p, B)
for _ in range(reps):
    eng.solve_device({k: v.data_ptr() for k, v in tin.items()}, {"grf": grf.data_ptr(), "status": st.data_ptr(), "iters": it.data_ptr()}, B)
pol = (it >> 16).double(); fac = (it & 0xFFFF).double()
print("kernel ms", eng.last_kernel_ms, "solves/s", B / eng.last_kernel_ms * 1e3, eng.last_stats(), "bad", int((st > 1).sum()),
      "polish rounds mean %.3f max %d" % (float(pol.mean()), int(pol.max())), "ipm iterations mean %.3f max %d" % (float((fac - pol).mean()), int((fac - pol).max())))
eng.close()
