#!/usr/bin/env python3
"""Host-buffer solve of a big batch: wall clock of the C ABI alone (Engine.solve with page-locked tensors) against the
facade Indy7SQP.solve (which also builds the stats dictionary) and the device-resident launch, for the pipeline
settings in the environment (I7_PIPE_CHUNK, I7_PIPE_STREAMS)."""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--horizon", type=int, default=32)
    ap.add_argument("--reps", type=int, default=20)
    args = ap.parse_args()
    import torch
    from indy7_mpc_b200 import Engine, Indy7SQP
    from oracle.model import load_model
    from tests.cases import make_cases
    B, N = args.batch, args.horizon
    xu, x0, goals = make_cases(load_model(), 32, N=N, seed=3)        # 32 seeded cases, tiled and jittered
    rng = np.random.default_rng(5)
    reps = (B + 31) // 32
    xu, x0, goals = (np.tile(a, (reps, 1))[:B].copy() for a in (xu, x0, goals))
    x0 += rng.normal(0, 1e-3, x0.shape)
    xu[:, :12] = x0
    eng = Engine(N=N, batch=B, dtype="fp64")
    d = lambda a: torch.tensor(a, device="cuda")
    out, _ = eng.solve(d(xu), d(x0), d(goals))            # a warm iterate
    xu = out.cpu().numpy()
    pin = lambda a: torch.tensor(a).pin_memory()
    p_xu, p_x0, p_g = pin(xu), pin(x0), pin(goals)
    d_xu, d_x0, d_g = d(xu), d(x0), d(goals)
    api = Indy7SQP(N=N, batch=B, dtype="fp64")

    def timed(f):
        ts = []
        for _ in range(args.reps + 3):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            f()
            torch.cuda.synchronize()
            ts.append((time.perf_counter() - t0) * 1e3)
        return float(np.median(ts[3:]))
    t_dev = timed(lambda: eng.solve(d_xu, d_x0, d_g))
    t_c = timed(lambda: eng.solve(p_xu, p_x0, p_g))
    t_np = timed(lambda: eng.solve(xu, x0, goals))
    t_api = timed(lambda: api.solve(p_x0, p_g, xu_init=p_xu))
    print(f"chunk={os.environ.get('I7_PIPE_CHUNK', '256')} streams={os.environ.get('I7_PIPE_STREAMS', '3')}: "
          f"device-resident {t_dev:.3f} ms | C ABI pinned {t_c:.3f} ms | C ABI pageable {t_np:.3f} ms | Indy7SQP.solve pinned {t_api:.3f} ms")


if __name__ == "__main__":
    main()
