"""indy7_mpc_b200 — B200-native SQP-MPC solve loop for the Indy7 arm (drop-in for the hot path of
A2R-Lab/indy7-mpc: src/osqp_solver.py + src/osqp_sqp.py).  CUDA only; importing the solver classes loads
libindy7_b200.so and fails loudly if it is missing."""
from .model import RobotModel, build_model_from_urdf, DEFAULT_URDF          # noqa: F401
from .utils import load_robot_model, rk4, print_stats                       # noqa: F401
from .engine import Engine                                                  # noqa: F401
from .osqp_solver import OSQPSolver                                         # noqa: F401
from .osqp_sqp import SQP_OSQP                                              # noqa: F401
from .osqp_mpc import MPC_OSQP                                              # noqa: F401
from .api import Indy7SQP                                                   # noqa: F401
from .tick import DeviceTick, ControllerTick                                # noqa: F401

__version__ = "0.1.0"
