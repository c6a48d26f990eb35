"""CPU oracle for the indy7-mpc SQP-MPC hot path.  TEST INFRASTRUCTURE ONLY.

This package is a plain numpy (fp64) restatement of the reference algorithm
(`src/osqp_solver.py`, `src/osqp_sqp.py`, `src/osqp_mpc.py:14-71`, `src/utils.py:3-18`) together with a
restatement of the published algorithms of the two third-party libraries that carry the arithmetic
and are NOT vendored in the reference:

  * Pinocchio (PyPI ``pin``, version unpinned by the reference) — URDF model semantics, forward
    kinematics, LOCAL_WORLD_ALIGNED joint Jacobian, ABA, ``computeABADerivatives`` (Featherstone,
    "Rigid Body Dynamics Algorithms", 2008; Carpentier & Mansard, RSS 2018).
  * OSQP (PyPI ``osqp``, version unpinned) — ADMM for ``min 1/2 z'Pz + q'z  s.t.  l <= Az <= u``
    (Stellato et al., Math. Prog. Comp. 2020), used by the reference with all-default settings.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import anything from here.  The product package ``indy7_mpc_b200`` never does: it fails
loudly when its CUDA library is missing instead of falling back to this code.

Parity pinning: the reference holds no unit tests; its stored notebook outputs are the
golden vectors.  ``tests/test_oracle_golden.py`` pins this oracle to them: 16 forward-kinematics
known-answers (to print precision) and the first values of the 500-step closed-loop goal-distance series of
``notebooks/pin_mpc_indy7.ipynb`` (steps 0..20 within 5e-6 relative with the exact-KKT QP back-end).
What stays UNPINNED: OSQP's raw iterate / iteration count (no vector in the reference records it); the
analytic derivatives and the Jacobian are pinned only jointly through the closed-loop series and are
self-validated here against complex-step differentiation.
"""
