"""safe_exploration_b200 -- B200-native (sm_100a) safe CEM-MPC + GP ellipsoid reachability.

Drop-in for the hot path of oscarkey/safe-exploration behind the reference's own interfaces:

    reference                                            here
    ---------------------------------------------------  ------------------------------------------
    ssm_cem.GpCemSSM (CemSSM)                            ssm.B200GpCemSSM
    gp_reachability_pytorch.onestep_reachability ...     reachability.onestep_reachability ...
    constrained_cem_mpc.ConstrainedCemMpc                cem.B200ConstrainedCemMpc
    safempc_cem.CemSafeMPC (SafeMPC)                     safempc.B200CemSafeMPC
    utils_config.create_solver (safempc_cem branch)      utils_config.create_solver
    episode_runner.do_rollout                            runner.do_rollout, runner.do_rollout_batch

All numerics run in libsecem.so (csrc/, C ABI in include/secem.h); there is no CPU fallback.
"""
from ._lib import SE_F32, SE_F64, SE_GEMM_AUTO, SE_GEMM_MMA, SE_GEMM_SIMT, SecemError, load as load_library
from .cem import (ActionConstraint, B200ConstrainedCemMpc, Constraint, DynamicsFunc, DynamicsFuncWrapper,
                  RolloutFunction, Rollouts, TorchPolytope, box2torchpoly, pick_rollout, sample_noise,
                  select_elites_refit)
from .reachability import (SafeProblem, is_ellipsoid_inside_polytope, lin_ellipsoid_safety_distance,
                           multistep_reachability, onestep_reachability)
from .safempc import (B200CemSafeMPC, B200SafeDynamics, EllipsoidObstacleConstraint, EllipsoidStateConstraint,
                      EllipsoidTerminalConstraint,
                      LqrFeedbackController, MpcResult, PQFlattener, SafeMPC, classify_objective,
                      construct_constraints, dlqr)
from .runner import do_rollout, do_rollout_batch
from .ssm import B200GpCemSSM, CemSSM, assert_shape
from .utils_config import create_cem_ssm, create_solver

__version__ = '0.1.0'
