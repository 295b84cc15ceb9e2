"""Quick device-timed headline number: python tools/dbg/qb.py [workload] [reps]"""
import sys, os, numpy as np, torch
sys.path.insert(0, os.getcwd())
import bench
from legged_mpc_control_b200 import abi
from legged_mpc_control_b200.engine import ConvexMpcEngine
wl = sys.argv[1] if len(sys.argv) > 1 else "headline"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 8
bench._WL = wl
B = 65536
p, batch = bench.make_workload(B, 0)
dev = torch.device("cuda", 0)
names = tuple(k for k in ("x0", "rot", "foot_pos_abs", "cmd", "contact", "q", "r", "mu") if k in batch)
tin = {k: torch.from_numpy(batch[k]).to(dev) for k in names}
W = abi.active_words(p.horizon)
tout = {"grf": torch.zeros(B, 12, dtype=torch.float64, device=dev), "status": torch.zeros(B, dtype=torch.int32, device=dev),
        "iters": torch.zeros(B, dtype=torch.int32, device=dev), "active_set": torch.zeros(B, W, dtype=torch.int32, device=dev)}
eng = ConvexMpcEngine(p, B, device=0)
pin = {k: v.data_ptr() for k, v in tin.items()}; pout = {k: v.data_ptr() for k, v in tout.items()}
flush = torch.empty(64 << 20, dtype=torch.int32, device=dev)
ms = []
for i in range(reps + 2):
    flush.fill_(i); torch.cuda.synchronize()
    eng.solve_device(pin, pout, B)
    if i >= 2: ms.append(eng.last_kernel_ms)
st = eng.last_stats()
print(f"QB {wl}: {B / (np.mean(ms) * 1e-3) / 1e6:.4f} M solves/s  kernel ms mean {np.mean(ms):.3f} min {np.min(ms):.3f}  fact {st['mean_factorizations']:.3f} status>1 {(tout['status'] > 1).sum().item()} status1 {(tout['status'] == 1).sum().item()} {os.environ.get('QB_TAG', '')}")
