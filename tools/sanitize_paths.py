"""Small factor+solve+matvec runs of every solve path, for compute-sanitizer (memcheck / racecheck / initcheck).
    compute-sanitizer --tool memcheck python tools/sanitize_paths.py
Prints one OK line per (path, shape); results are compared with the oracle so a corrupted run is noticed."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from block_diag_ilu_b200 import _lib, device as dev, synth  # noqa: E402
from oracle import loader  # noqa: E402

PATHS = {"auto": 0, "chain": 1, "wave_generic": 2, "wave32_onerow": 3, "mrhs": 4, "mrhs_wave": 6, "skew32": 8, "pf": 9, "half32": 10}
SHAPES = [(2, 130, 32, 2, 1), (1, 70, 64, 3, 16), (20, 129, 8, 2, 1), (1, 200, 16, 1, 1), (3, 40, 4, 1, 2)]
only = sys.argv[1:] or list(PATHS)
orc = loader.best()
L = _lib.lib()
for name in only:
    old = L.bdilu_debug_force_chain_solve(PATHS[name])
    for (nsys, N, w, d, nrhs) in SHAPES:
        data, b = synth.make_batch(nsys, N, w, d, "D", seed0=N, nrhs=nrhs)
        dd, bb = torch.from_numpy(data).cuda(), torch.from_numpy(b).cuda()
        ilu = dev.BatchedILU(dd, N, w, d, copy=True)
        x = ilu.solve(bb)
        r2 = dev.residual_norm2(dd, bb, x, N, w, d)
        torch.cuda.synchronize()
        ref = data[0].copy()
        _, rbc, cbr, _ = orc.factor(ref, N, w, d, 0, w)
        xr, _ = orc.solve(ref, N, w, d, 0, w, rbc, cbr, b[0, 0])
        err = np.abs(x[0, 0].cpu().numpy() - xr).max() / np.abs(xr).max()
        assert err < 1e-12 and torch.isfinite(r2).all(), (name, N, w, err)
        print("OK", name, (nsys, N, w, d, nrhs), "err %.1e" % err, flush=True)
    L.bdilu_debug_force_chain_solve(old)
