import sys, os, time, ctypes as C, numpy as np
sys.path.insert(0, os.getcwd())
from legged_mpc_control_b200 import presets, synth, _lib
from legged_mpc_control_b200.engine import ShardedConvexMpcEngine, ConvexMpcEngine
import torch
ndev = torch.cuda.device_count()
p = presets.preset("northstar_a1")
L = _lib.lib()
def pinned_copy(a):
    ptr = L.lmpc_alloc_pinned(a.nbytes)
    out = np.frombuffer((C.c_char * a.nbytes).from_address(ptr), dtype=a.dtype).reshape(a.shape)
    out[...] = a
    return out
for G in range(1, ndev + 1):
    B = 65536 * G
    b = synth.make_batch(B, p.horizon, p.dt, seed=3)
    for mode in ("torch_pin", "lmpc_pin", "pageable"):
        names = ("x0", "rot", "foot_pos_abs", "cmd", "contact")
        if mode == "torch_pin": hb = {k: torch.from_numpy(b[k]).pin_memory().numpy() for k in names}
        elif mode == "lmpc_pin": hb = {k: pinned_copy(b[k]) for k in names}
        else: hb = {k: b[k] for k in names}
        sh = ShardedConvexMpcEngine(p, B, list(range(G)))
        for _ in range(2): sh.compute_grfs(hb, copy=False)
        t0 = time.perf_counter()
        for _ in range(5): o = sh.compute_grfs(hb, copy=False)
        dt = (time.perf_counter() - t0) / 5
        print(f"G={G} {mode}: {B / dt / 1e6:.3f} M solves/s  {dt * 1e3:.2f} ms per call, kernel ms per shard", [L.lmpc_last_kernel_ms(L.lmpc_sharded_handle(sh._s, g)) for g in range(G)])
        sh.close()
