"""Constrained CEM-MPC on the B200 behind the reference's `ConstrainedCemMpc` interface
(constrained-cem-mpc/constrained_cem_mpc/constrained_cem_mpc.py:195-280).

The whole optimisation -- sampling, T ellipsoid steps per candidate, constraint costs, elite selection, refit, final
pick -- runs on the device inside one asynchronous `se_cem_solve` call; this module only owns tensors and mirrors the
reference's plug-in types (`TorchPolytope`, `Constraint`s, `DynamicsFunc`, `Rollouts`).
"""
import ctypes as C
from abc import ABC, abstractmethod
from collections import namedtuple
from typing import List, Optional, Sequence, Tuple, Union

import numpy as np
import torch
from torch import Tensor

from . import _lib
from .reachability import SafeProblem, check_flags
from .ssm import B200GpCemSSM, assert_shape

# constrained_cem_mpc.py:127
Rollouts = namedtuple('Rollouts', 'trajectories actions objective_costs constraint_costs')


class TorchPolytope:
    """{x : A x <= b} (constrained_cem_mpc.py:13-53).  Accepts anything with `.A`, `.b` (e.g. polytope.Polytope) or a
    pair (A, b)."""

    def __init__(self, polytope):
        A, b = (polytope.A, polytope.b) if hasattr(polytope, 'A') else polytope
        self._A = torch.as_tensor(np.asarray(A, dtype=np.float64))
        self._b = torch.as_tensor(np.asarray(b, dtype=np.float64)).reshape(-1)
        self.dim = self._A.shape[1]
        cheb = getattr(polytope, 'chebXc', None)
        self.chebXc = None if cheb is None else torch.as_tensor(np.asarray(cheb, dtype=np.float64))

    def __contains__(self, x: Tensor) -> bool:
        assert_shape(x, (self.dim,))
        return not bool((self._A.to(x) @ x - self._b.to(x) > 0).any())

    def contains_points(self, xs: Tensor) -> int:
        """Number of rows of xs [N, dim] inside the polytope (boundary counts as inside: strict `>`)."""
        assert xs.dim() == 2 and xs.shape[1] == self.dim, f'Wanted (N, {self.dim}) got {xs.shape}'
        outside = ((xs @ self._A.to(xs).t() - self._b.to(xs)) > 0).any(dim=1)
        return int(xs.shape[0] - outside.sum())

    def to(self, device):
        self._A = self._A.to(device)
        self._b = self._b.to(device)
        if self.chebXc is not None:
            self.chebXc = self.chebXc.to(device)
        return self


def box2torchpoly(box) -> TorchPolytope:
    """polytope.box2poly semantics (A = [I; -I], b = [hi; -lo]) -> TorchPolytope (constrained_cem_mpc.py:57-59)."""
    box = np.asarray(box, dtype=np.float64)
    n = box.shape[0]
    poly = TorchPolytope((np.vstack((np.eye(n), -np.eye(n))), np.concatenate((box[:, 1], -box[:, 0]))))
    poly.chebXc = torch.as_tensor(box.mean(axis=1))
    return poly


class Constraint(ABC):
    """constrained_cem_mpc.py:62-68: cost >= 0 of one trajectory, 0 iff feasible."""

    @abstractmethod
    def __call__(self, trajectory: Tensor, actions: Tensor) -> float:
        pass


class ActionConstraint(Constraint):
    """penalty x #{t : u_t outside the polytope} (constrained_cem_mpc.py:100-110).  Evaluated on the device for every
    rollout; calling the object evaluates one rollout on the host with the same rule."""

    def __init__(self, safe_region: TorchPolytope, penalty: float = 3):
        self._self_region = safe_region
        self._penalty = penalty

    def __call__(self, trajectory: Tensor, actions: Tensor) -> float:
        return self._penalty * (actions.shape[0] - self._self_region.contains_points(actions))

    def batch_call(self, trajectories: Tensor, actions: Tensor) -> Tensor:
        """The same rule for all rollouts at once: actions [N, T, n_u] -> costs [N] (device tensor ops)."""
        region = self._self_region
        viol = (torch.matmul(actions, region._A.to(actions).t()) - region._b.to(actions)) > 0
        return self._penalty * viol.any(dim=2).sum(dim=1).to(actions.dtype)


class DynamicsFunc(ABC):
    """constrained_cem_mpc.py:113-124."""

    @abstractmethod
    def __call__(self, states: Tensor, actions: Tensor) -> Tuple[Tensor, Tensor]:
        pass


class DynamicsFuncWrapper(DynamicsFunc):
    """Wraps a function as a DynamicsFunc (safempc_cem.py:149-156)."""

    def __init__(self, func):
        self._func = func

    def __call__(self, states: Tensor, actions: Tensor) -> Tuple[Tensor, Tensor]:
        return self._func(states, actions)


class RolloutFunction:
    """The reference's RolloutFunction (constrained_cem_mpc.py:130-192) on DEVICE tensors: the interface path of
    `B200ConstrainedCemMpc` for dynamics / constraints that are arbitrary Python objects.  The candidates
    are drawn by the library's sampler (`se_cem_noise`, Philox keyed by seed / iteration / rollout id), the dynamics
    callback is called T times with [N, S] device tensors exactly as in the reference, constraints that know how to
    evaluate a whole batch (`batch_call`) do so on the device and any other `Constraint` is called per rollout like the
    reference does (:187-192)."""

    def __init__(self, dynamics: DynamicsFunc, constraints: Sequence[Constraint], state_dimen: int, action_dimen: int,
                 time_horizon: int, num_rollouts: int, seed: int = 0):
        self._dynamics = dynamics
        self._constraints = list(constraints)
        self._state_dimen = state_dimen
        self._action_dimen = action_dimen
        self._time_horizon = time_horizon
        self._num_rollouts = num_rollouts
        self.seed = int(seed)
        self.iteration = 0   # set by the optimiser before every call (the sampler's counter)
        self.noise = None    # optional RNG-fixed standard-normal noise [N, T, n_u] of the next call

    def perform_rollouts(self, args: Tuple[Tensor, Tensor, Tensor]) -> Rollouts:
        """(initial_state [S], means [T, n_u], stds [T, n_u]) -> Rollouts (:145-156)."""
        initial_state, means, stds = args
        _lib.require_cuda(initial_state, 'initial_state')
        initial_states = initial_state.reshape(1, -1).repeat((self._num_rollouts, 1))
        trajectories, actions, objective_costs = self._sample_trajectory(initial_states, means, stds)
        constraint_costs = self._compute_constraint_costs(trajectories, actions)
        return Rollouts(trajectories, actions, objective_costs, constraint_costs)

    def _sample_trajectory(self, initial_states: Tensor, means: Tensor, stds: Tensor):
        N, T, n_u, S = self._num_rollouts, self._time_horizon, self._action_dimen, self._state_dimen
        assert_shape(initial_states, (N, S))
        assert_shape(means, (T, n_u))
        assert_shape(stds, (T, n_u))
        dt, dev = initial_states.dtype, initial_states.device
        eps = self.noise if self.noise is not None else \
            sample_noise(self.seed, self.iteration, 1, N, 0, N, T, n_u, dtype=dt, device=dev)[0]
        assert_shape(eps, (N, T, n_u))
        # Normal(means, stds).sample((N,)) == means + stds * eps  (:167)
        actions = means.to(dt).unsqueeze(0) + stds.to(dt).unsqueeze(0) * eps.to(device=dev, dtype=dt)
        rollout_batch = getattr(self._dynamics, 'rollout_batch', None)
        if rollout_batch is not None:
            out = rollout_batch(initial_states, actions)  # all T steps in one device call when the dynamics can
            if out is not None:
                return out[0], actions, out[1]
        trajectories = torch.empty((N, T + 1, S), dtype=dt, device=dev)
        trajectories[:, 0, :] = initial_states
        objective_costs = torch.zeros((N,), dtype=dt, device=dev)
        for t in range(T):
            next_states, costs = self._dynamics(trajectories[:, t, :], actions[:, t, :])
            assert_shape(next_states, (N, S))
            assert_shape(costs, (N,))
            trajectories[:, t + 1, :] = next_states
            objective_costs += costs.to(dt)
        return trajectories, actions, objective_costs

    def _compute_constraint_costs(self, trajectories: Tensor, actions: Tensor) -> Tensor:
        N = trajectories.shape[0]
        costs = torch.zeros((N,), dtype=trajectories.dtype, device=trajectories.device)
        for constraint in self._constraints:
            batch = getattr(constraint, 'batch_call', None)
            if batch is not None:
                costs += batch(trajectories, actions).to(costs)
            else:  # an arbitrary Constraint: one call per rollout, as the reference does (:187-192)
                costs += torch.as_tensor([float(constraint(trajectories[i], actions[i])) for i in range(N)],
                                         dtype=costs.dtype, device=costs.device)
        return costs


class B200ConstrainedCemMpc:
    """Drop-in for `ConstrainedCemMpc` (constrained_cem_mpc.py:195-280).  Construction mirrors the reference (:204-220).

    Two evaluation paths behind the same methods:
      * fused -- `dynamics_func` is a `safempc.B200SafeDynamics` on a `B200GpCemSSM` with a built-in objective, and every
        constraint is one of ActionConstraint / Ellipsoid{Terminal,State,Obstacle}Constraint: dynamics and constraints
        are folded into the device-side problem and the whole optimisation is one `se_cem_solve`.  `state_dimen` is
        then the flat dimension n_s + n_s^2 (PQFlattener).
      * interface -- anything else (an arbitrary `DynamicsFunc`, an unknown `Constraint`, a CemSSM that is not the B200
        GP, an injected `rollout_function`): the reference-shaped loop of `optimize_trajectories` runs on device tensors
        with the Python callbacks called exactly as the reference calls them, and everything that is not a callback stays
        on the library's kernels (sampling `se_cem_noise`, elite selection + refit `se_cem_select_refit`, final pick
        `se_cem_pick`).  It is an interface fallback, not a CPU fallback: tensors must live on a CUDA device.

    Extras over the reference: `noise` (RNG-fixed candidates), `seed`, `pick_best`, and on the fused path `record`,
    batched instances and population sharding over ranks (`set_comm`).
    """

    def __init__(self, dynamics_func, constraints: Sequence[Constraint], state_dimen: int, action_dimen: int,
                 time_horizon: int, num_rollouts: int, num_elites: int, num_iterations: int, rollout_function=None,
                 seed: int = 0, pick_best: bool = False, record: str = 'full', num_instances: int = 1,
                 rank: int = 0, world_size: int = 1, first_instance: int = 0):
        from .safempc import B200SafeDynamics, can_fold, fold_constraints  # local import: safempc imports this module
        if record not in ('full', 'costs', 'none'):
            raise ValueError(f'record must be full|costs|none, got {record}')
        if num_rollouts % world_size != 0:
            raise ValueError(f'num_rollouts {num_rollouts} must be divisible by world_size {world_size}')
        self._state_dimen = state_dimen
        self._action_dimen = action_dimen
        self._time_horizon = time_horizon
        self._num_rollouts = num_rollouts
        self._num_elites = num_elites
        self._num_iterations = num_iterations
        self._constraints = list(constraints) if constraints is not None else []
        self.seed = int(seed)
        self.fused = (rollout_function is None and isinstance(dynamics_func, B200SafeDynamics) and
                      dynamics_func.fusable and all(can_fold(c) for c in self._constraints))
        self._dyn = dynamics_func
        if self.fused:
            self._ssm: B200GpCemSSM = dynamics_func.ssm
            n_s = self._ssm.num_states
            if state_dimen != n_s + n_s * n_s:
                raise ValueError(f'Wanted state_dimen {n_s + n_s * n_s}, got {state_dimen}')
            if action_dimen != self._ssm.num_actions:
                raise ValueError(f'Wanted action_dimen {self._ssm.num_actions}, got {action_dimen}')
            self._rollout_function = self  # safempc_cem.py:270 reads mpc._rollout_function._constraints
            self._problem: SafeProblem = fold_constraints(dynamics_func.problem, self._constraints)
        else:
            if num_instances != 1 or world_size != 1:
                raise TypeError('batched instances and population sharding need the fused path (B200SafeDynamics on a '
                                'B200GpCemSSM with built-in objective and constraints)')
            self._ssm = getattr(dynamics_func, 'ssm', None)
            if rollout_function is None:
                rollout_function = RolloutFunction(dynamics_func, self._constraints, state_dimen, action_dimen,
                                                   time_horizon, num_rollouts, seed=self.seed)
            self._rollout_function = rollout_function
            self._problem = None
        self.pick_best = bool(pick_best)
        self.record = record
        self.num_instances = int(num_instances)
        self.first_instance = int(first_instance)  # instance sharding: global index of this solver's instance 0
        self.rank, self.world_size = int(rank), int(world_size)
        self._lib = _lib.load()
        self._handle = C.c_void_p(None)
        self._handle_key = None
        self._comm = None
        self.last_info = {}
        self._out_cache = {}  # persistent result tensors of solve(reuse_outputs=True)

    # ---- handle -------------------------------------------------------------------------------------------------
    def _destroy(self):
        if getattr(self, '_handle', None):
            self._lib.se_cem_destroy(self._handle)
            self._handle = C.c_void_p(None)
            self._handle_key = None

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    def set_objective(self, objective: int, objective_theta: float = 0.0) -> None:
        """Change the built-in objective selector / its target (the InvertedPendulum advances its target angle during
        an episode, environments.py:488-515); the device-side problem is rebuilt on the next solve."""
        if not self.fused:
            raise _lib.SecemError('set_objective applies to the fused path')
        if (int(objective), float(objective_theta)) != (self._problem.objective, self._problem.objective_theta):
            self._problem = self._problem.with_(objective=int(objective), objective_theta=float(objective_theta))
            self._dyn.problem = self._dyn.problem.with_(objective=int(objective), objective_theta=float(objective_theta))
            self._destroy()

    def set_comm(self, nccl_comm: Optional[int], rank: int, world_size: int) -> None:
        """Attach an NCCL communicator made by `dist.create_nccl_comm` (population sharding)."""
        if not self.fused and world_size != 1:
            raise TypeError('population sharding needs the fused path')
        if self._num_rollouts % world_size != 0:
            raise ValueError(f'num_rollouts {self._num_rollouts} must be divisible by world_size {world_size}')
        self._comm, self.rank, self.world_size = nccl_comm, int(rank), int(world_size)
        self._destroy()

    def _ensure_handle(self):
        if not self.fused:
            raise _lib.SecemError('this solver runs the interface path (Python dynamics / constraints): it has no '
                                  'device-side CEM handle; use get_actions / optimize_trajectories')
        gp_handle = self._ssm.handle
        # never key on the handle's address: update_model destroys and re-creates the model and the allocator readily
        # returns the same address (the library checks the generation too and refuses a stale GP handle)
        key = (self._ssm.generation, self.num_instances, self.pick_best, self.rank, self.world_size,
               self.first_instance)
        if self._handle and self._handle_key == key:
            return
        self._destroy()
        n_local = self._num_rollouts // self.world_size
        cfg = _lib.SeCemConfig(self.num_instances, self._num_rollouts, self._time_horizon, self._num_elites,
                               self._num_iterations, 1 if self.pick_best else 0, self.rank * n_local, n_local,
                               self.first_instance)
        cs = self._problem.c_struct()
        handle = C.c_void_p(None)
        with torch.cuda.device(self._ssm.device):
            _lib.check(self._lib.se_cem_create(C.byref(handle), gp_handle, C.byref(cs), C.byref(cfg)))
            if self.world_size > 1:
                _lib.check(self._lib.se_cem_set_comm(handle, C.c_void_p(self._comm), self.rank, self.world_size))
        self._handle, self._handle_key = handle, key

    # ---- solve ----------------------------------------------------------------------------------------------------
    def solve(self, p0: Tensor, noise: Optional[Tensor] = None, seed: Optional[int] = None, sync: bool = True,
              reuse_outputs: bool = False):
        """Run the CEM optimisation for M = p0.shape[0] instances.

        :param p0: [M, n_s] initial point states on the device
        :param noise: None (on-device Philox) or [I, M, N, T, n_u] standard-normal noise
        :param reuse_outputs: write into this object's persistent result tensors (and stage p0 in a persistent
               buffer) instead of fresh ones: the returned tensors are overwritten by the next such call, and the
               library sees the same buffers every time, so it replays its captured CUDA graph of the solve
        :returns: dict of device tensors: actions [M, T, n_u], found [M], best_id [M], means / stds [I, M, T, n_u],
                  elite_ids [I, M, E], and per `record`: objective / constraint [I, M, N_local],
                  rollout_actions [I, M, N_local, T, n_u], trajectories [I, M, N_local, T+1, S]
        """
        ssm = self._ssm
        n_s, n_u, T, I, E = ssm.num_states, ssm.num_actions, self._time_horizon, self._num_iterations, self._num_elites
        M = p0.shape[0]
        assert_shape(p0, (M, n_s))
        if M != self.num_instances:
            self.num_instances = M
        self._ensure_handle()
        dev, dt = ssm.device, ssm.dtype
        p0 = ssm._prep(p0, 'state')
        n_local = self._num_rollouts // self.world_size
        S = n_s + n_s * n_s
        new = lambda shape, dtype=dt: torch.empty(shape, dtype=dtype, device=dev)
        cache_key = (M, self.record, dt, str(dev), n_local)
        cached = self._out_cache.get(cache_key) if reuse_outputs else None
        if cached is not None:
            res, p0_buf = cached
            if p0.data_ptr() != p0_buf.data_ptr():
                p0_buf.copy_(p0)
            p0 = p0_buf
        else:
            # actions | found | best_id | flags share one allocation, so a caller that only wants the outcome of the
            # solve (get_actions_point) reads it back with a single device -> host copy
            esz = torch.empty((), dtype=dt).element_size()
            a_bytes = M * T * n_u * esz
            pack = torch.zeros(a_bytes + 8 * M + 16, dtype=torch.uint8, device=dev)
            res = dict(actions=pack[:a_bytes].view(dt).view(M, T, n_u),
                       found=pack[a_bytes:a_bytes + 4 * M].view(torch.int32),
                       best_id=pack[a_bytes + 4 * M:a_bytes + 8 * M].view(torch.int32),
                       means=new((I, M, T, n_u)), stds=new((I, M, T, n_u)), elite_ids=new((I, M, E), torch.int32),
                       flags=pack[a_bytes + 8 * M:].view(torch.int32), _pack=pack)
            if self.record in ('full', 'costs'):
                res['objective'] = new((I, M, n_local))
                res['constraint'] = new((I, M, n_local))
            if self.record == 'full':
                res['rollout_actions'] = new((I, M, n_local, T, n_u))
                res['trajectories'] = new((I, M, n_local, T + 1, S))
            if reuse_outputs:
                p0_buf = p0.clone()
                self._out_cache = {cache_key: (res, p0_buf)}  # one live set: a new shape replaces the old buffers
                p0 = p0_buf
        if noise is not None:
            assert_shape(noise, (I, M, self._num_rollouts, T, n_u))
            noise = ssm._prep(noise, 'noise')
        out = _lib.SeCemOutputs()
        for name, _ in _lib.SeCemOutputs._fields_:
            t = res.get(name)
            setattr(out, name, None if t is None else t.data_ptr())
        if seed is None:
            seed = self.seed
            self.seed += 1  # a fresh candidate stream per solve, like successive draws from torch's generator
        with torch.cuda.device(dev):
            _lib.check(self._lib.se_cem_solve(self._handle, _lib.ptr(p0), _lib.ptr(noise), C.c_uint64(int(seed)),
                                              C.byref(out), _lib.stream_ptr(dev)))
        if sync:
            check_flags(self._collective_flags(res['flags']), 'CEM rollouts', dump=(self._ssm, p0, None))
        return res

    def _collective_flags(self, flags: Tensor) -> Tensor:
        """Host copy of the numeric flags.  Sharded: the element-wise maximum over the ranks, so that every rank raises
        (or warns) together instead of one rank leaving the others blocked in the next all-gather."""
        if self.world_size > 1:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                f = flags.clone() if dist.get_backend() == 'nccl' else flags.cpu()
                dist.all_reduce(f, op=dist.ReduceOp.MAX)
                return f.cpu()
        return flags.cpu()

    def get_actions_point(self, state) -> Optional['np.ndarray']:
        """What CemSafeMPC.get_action needs from get_actions (safempc_cem.py:236-241) for a POINT state given on the
        host: the action sequence [T, n_u] of a feasible rollout of the last iteration as a NumPy array, or None.
        One pinned host -> device copy of the state, one device -> host copy of (actions, found, flags), one stream
        synchronisation; no per-iteration records are exposed (use get_actions for those)."""
        import numpy as np
        ssm = self._ssm
        if not self.fused:
            n_s = len(np.asarray(state).reshape(-1))
            dt = getattr(ssm, 'dtype', torch.float64)
            dev = getattr(ssm, 'device', torch.device('cuda:0'))
            flat = torch.zeros(self._state_dimen, dtype=dt, device=dev)
            flat[:n_s] = torch.as_tensor(np.asarray(state, dtype=np.float64).reshape(-1), dtype=dt)
            actions, _ = self.get_actions(flat)
            return None if actions is None else actions.detach().cpu().numpy()
        n_s, T, n_u = ssm.num_states, self._time_horizon, ssm.num_actions
        dev, dt = ssm.device, ssm.dtype
        n_local = self._num_rollouts // self.world_size
        cached = self._out_cache.get((1, self.record, dt, str(dev), n_local))
        stage = getattr(self, '_point_stage', None)
        if cached is None or stage is None or stage[2] is not cached[0]['_pack']:
            # first call (or new buffers): the general path creates the persistent buffers
            p0 = torch.as_tensor(np.asarray(state, dtype=np.float64).reshape(1, n_s), dtype=dt, device=dev)
            res = self.solve(p0, sync=True, reuse_outputs=True)
            self.last_info = res
            pack = res['_pack']
            self._point_stage = (torch.empty((1, n_s), dtype=dt).pin_memory(),
                                 torch.empty(pack.numel(), dtype=torch.uint8).pin_memory(), pack)
            return res['actions'][0].cpu().numpy() if int(res['found'][0]) == 1 else None
        res, p0_buf = cached
        h_p0, h_pack, pack = stage
        h_p0.numpy()[0, :] = np.asarray(state, dtype=np.float64).reshape(n_s)
        p0_buf.copy_(h_p0, non_blocking=True)
        res = self.solve(p0_buf, sync=False, reuse_outputs=True)
        self.last_info = res
        h_pack.copy_(pack, non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()
        a_bytes = T * n_u * h_p0.element_size()
        tail = h_pack[a_bytes:].view(torch.int32)  # found, best_id, flags[4]
        check_flags(tail[2:], 'CEM rollouts', dump=(self._ssm, p0_buf, None))
        if int(tail[0]) != 1:
            return None
        return h_pack[:a_bytes].view(dt).view(T, n_u).numpy().copy()

    def _flat_state_to_p0(self, state: Tensor) -> Tensor:
        n_s = self._ssm.num_states
        state = state.reshape(1, -1) if state.dim() == 1 else state
        assert_shape(state, (state.shape[0], self._state_dimen))
        if bool((state[:, n_s:] != 0).any()):
            raise NotImplementedError('the fused CEM path starts from point states (q = 0), as CemSafeMPC.get_action does')
        return state[:, :n_s]

    def _rollouts_of(self, res, inst: int = 0) -> List[Rollouts]:
        out = []
        for i in range(self._num_iterations):
            out.append(Rollouts(res['trajectories'][i, inst] if 'trajectories' in res else None,
                                res['rollout_actions'][i, inst] if 'rollout_actions' in res else None,
                                res['objective'][i, inst] if 'objective' in res else None,
                                res['constraint'][i, inst] if 'constraint' in res else None))
        return out

    def _optimize_interface(self, initial_state: Tensor, noise: Optional[Tensor]) -> List[Rollouts]:
        """optimize_trajectories (constrained_cem_mpc.py:222-239) on the interface path: Python callbacks for dynamics /
        constraints, library kernels for sampling, elite selection (+ index tie-break) and the refit."""
        _lib.require_cuda(initial_state, 'initial_state')
        rf = self._rollout_function
        dev = initial_state.device
        dt = initial_state.dtype if initial_state.is_floating_point() else torch.get_default_dtype()
        means = torch.zeros((self._time_horizon, self._action_dimen), dtype=dt, device=dev)
        stds = torch.ones((self._time_horizon, self._action_dimen), dtype=dt, device=dev)
        seed = self.seed
        self.seed += 1
        rollouts_by_time = []
        self.last_info = dict(means=[], stds=[], elite_ids=[])
        for i in range(self._num_iterations):
            if isinstance(rf, RolloutFunction):
                rf.seed, rf.iteration = seed, i
                rf.noise = None if noise is None else noise[i]
            rollouts = rf.perform_rollouts((initial_state, means, stds))
            ids, means, stds = select_elites_refit(rollouts.objective_costs, rollouts.constraint_costs,
                                                   rollouts.actions, self._num_elites)
            rollouts_by_time.append(rollouts)
            self.last_info['means'].append(means), self.last_info['stds'].append(stds)
            self.last_info['elite_ids'].append(ids)
        return rollouts_by_time

    def optimize_trajectories(self, initial_state: Tensor, noise: Optional[Tensor] = None) -> List[Rollouts]:
        """constrained_cem_mpc.py:222-239 -> rollouts of every iteration (final one last)."""
        if not self.fused:
            return self._optimize_interface(initial_state, noise)
        res = self.solve(self._flat_state_to_p0(initial_state), None if noise is None else noise.unsqueeze(1))
        self.last_info = res
        return self._rollouts_of(res)

    def get_actions(self, state: Tensor, noise: Optional[Tensor] = None, reuse_outputs: bool = False
                    ) -> Tuple[Union[Tensor, None], List[Rollouts]]:
        """constrained_cem_mpc.py:259-280: (actions [T, n_u] of a feasible rollout of the last iteration | None,
        rollouts by iteration).  `noise` [I, N, T, n_u] fixes the candidates (parity tests)."""
        if not self.fused:
            _lib.require_cuda(state, 'state')
            rollouts_by_time = self._optimize_interface(state, noise)
            last = rollouts_by_time[-1]
            best = pick_rollout(last.objective_costs, last.constraint_costs, self.pick_best)
            self.last_info['best_id'] = best
            return (last.actions[best] if best >= 0 else None), rollouts_by_time
        res = self.solve(self._flat_state_to_p0(state), None if noise is None else noise.unsqueeze(1),
                         reuse_outputs=reuse_outputs)
        self.last_info = res
        actions = res['actions'][0] if int(res['found'][0]) == 1 else None
        return actions, self._rollouts_of(res)

    def get_actions_batch(self, states: Tensor, noise: Optional[Tensor] = None, reuse_outputs: bool = False):
        """M independent MPC problems side by side (BASELINE config "batched independent MPC solves").
        :param states: [M, n_s] point states -> actions [M, T, n_u], found [M] (bool)"""
        res = self.solve(states, noise, reuse_outputs=reuse_outputs)
        self.last_info = res
        return res['actions'], res['found'].bool()


def select_elites_refit(objective_costs: Tensor, constraint_costs: Tensor, actions: Tensor, num_elites: int):
    """ConstrainedCemMpc._select_elites + the refit of optimize_trajectories (constrained_cem_mpc.py:234-257) on given
    costs, evaluated by the same device kernel the solver uses.  -> elite_ids [E] (int32), mean [T, n_u], std [T, n_u]"""
    _lib.require_cuda(objective_costs, 'objective_costs')
    lib = _lib.load()
    N, T, n_u = actions.shape
    dt, dev = actions.dtype, actions.device
    if dt not in (torch.float32, torch.float64):
        dt = torch.float64
    code = _lib.dtype_code(dt)
    obj = objective_costs.detach().to(dt).contiguous()
    con = constraint_costs.detach().to(dt).contiguous()
    act = actions.detach().to(dt).contiguous()
    ids = torch.empty((num_elites,), dtype=torch.int32, device=dev)
    mean = torch.empty((T, n_u), dtype=dt, device=dev)
    std = torch.empty((T, n_u), dtype=dt, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.se_cem_select_refit(code, _lib.ptr(obj), _lib.ptr(con), _lib.ptr(act), N, T, n_u, num_elites,
                                           _lib.ptr(ids), _lib.ptr(mean), _lib.ptr(std), _lib.stream_ptr(dev)))
    return ids, mean, std


def pick_rollout(objective_costs: Tensor, constraint_costs: Tensor, pick_best: bool = False) -> int:
    """The rollout `get_actions` returns (constrained_cem_mpc.py:272-280) on given costs, by the device kernel the solver
    uses (`se_cem_pick`): the first feasible one (quirk Q3), or the arg-min objective feasible one; -1 if none."""
    _lib.require_cuda(constraint_costs, 'constraint_costs')
    lib = _lib.load()
    dt = objective_costs.dtype if objective_costs.is_floating_point() else torch.float64
    obj = objective_costs.detach().to(dt).contiguous()
    con = constraint_costs.detach().to(dt).contiguous()
    best = torch.empty((1,), dtype=torch.int32, device=con.device)
    with torch.cuda.device(con.device):
        _lib.check(lib.se_cem_pick(_lib.dtype_code(dt), _lib.ptr(obj), _lib.ptr(con), con.numel(),
                                   1 if pick_best else 0, _lib.ptr(best), _lib.stream_ptr(con.device)))
    return int(best.item())


def sample_noise(seed: int, iteration: int, num_instances: int, num_rollouts: int, first: int, count: int, T: int,
                 n_u: int, dtype=torch.float64, device='cuda:0') -> Tensor:
    """The candidate noise `solve` draws on the device (Philox4x32-10) -> [M, count, T, n_u]."""
    lib = _lib.load()
    out = torch.empty((num_instances, count, T, n_u), dtype=dtype, device=device)
    with torch.cuda.device(out.device):
        _lib.check(lib.se_cem_noise(_lib.dtype_code(dtype), C.c_uint64(seed), iteration, num_instances, num_rollouts,
                                    first, count, T, n_u, _lib.ptr(out), _lib.stream_ptr(out.device)))
    return out
