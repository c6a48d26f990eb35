from indy7_mpc_b200.batch_sqp import *          # noqa: F401,F403
from indy7_mpc_b200.batch_sqp import SUPPORTED_BATCH  # noqa: F401
