"""robotic-mpc_b200 -- B200-native batched engine for the UR10 surface-tracking MPC hot path of lynet55/robotic-mpc.

The directory name contains a hyphen (it is the repository's package name); import it through the
``robotic_mpc_b200`` symlink at the repository root or with ``importlib.import_module("robotic-mpc_b200")``.
"""
from .params import RK, RL, RP, RS, pack_params  # noqa: F401
from .engine import SQP, SQP_RTI, Engine, RmpcError, load_library, solve_static  # noqa: F401
from .surface import Surface  # noqa: F401,E402
from .prediction_model import SixDofRobot, UrdfLoader  # noqa: F401,E402
from .simulation_model import Robot  # noqa: F401,E402
from .trajectory_optimizer import MPC, BatchedSolver, SolverOptions, SolverView  # noqa: F401,E402
from .simulator import SimulationManager, Simulator, run_batch, save_result_npz, summary_from_row  # noqa: F401,E402
from . import robots, terminal_constraint  # noqa: F401,E402
