import os, sys, numpy as np, torch
sys.path.insert(0, os.getcwd())
from spmv_b200 import api
ex = api.B200Executor.create(0); ex.use_torch_stream()
for name in ["lap27_6x5x8", "lap7_12x12x12", "powerlaw_3000"]:
  for sym in (False, True):
    g = np.load(f"tests/golden/cg_{name}.npz"); n = len(g["b"])
    A = api.Matrix.create_matrix(ex, g["rowptr"], g["colind"], g["values"], n, n, [], [], sym, api.P2P_BLOCKING)
    b = torch.from_numpy(g["b"]).cuda(); x = torch.zeros(n, dtype=torch.float64, device="cuda")
    it, hist = api.cg(ex, A, b, x, int(g["kmax"]), float(g["rtol"]), history=True)
    w = g["history"]; m = min(it, len(w))
    d = np.abs(hist[:m]-w[:m])/w[:m]
    print(name, sym, it, int(g["iterations"]), "dev first5", d[:5], "max", d.max(), "argmax", d.argmax(), "hist0", hist[:3], w[:3])
    A.close()
