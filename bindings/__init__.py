"""``bindings`` package so that the reference drivers' ``import bindings.batch_sqp`` /
``from bindings.batch_sqp import ...`` (gato_controller.py:8, src/gato_mpc_batch.py:5) resolve to the B200 backend."""
