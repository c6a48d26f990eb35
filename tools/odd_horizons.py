import sys, numpy as np
sys.path.insert(0, '/root/repo')
from indy7_mpc_b200 import Engine
from oracle.cport import COracle
from oracle.model import load_model
from tests.cases import make_cases
om = load_model()
for N in (2, 3, 4, 7, 33, 47, 100):
    xu, x0, goals = make_cases(om, 3, N=N, seed=N)
    orc = COracle(om, N=N, backend="kkt", batch=3)
    want, st_o, _ = orc.sqp(x0, goals, xu)
    for cl in (1, 8, 4, 2):
        try:
            eng = Engine(N=N, batch=3, cluster=cl)
        except Exception as e:
            print(N, cl, "create failed:", e); continue
        out, st = eng.solve(xu, x0, goals)
        out = np.asarray(out); err = np.abs(out - want).max() / np.abs(want).max()
        print(N, cl, eng.kernel_form(), f"rel err {err:.2e}", "alphas", np.asarray(st)[:, 4:6].tolist() == st_o[:, 4:6].tolist())
        eng.close()
