"""Episode rollouts around the solver: the control loop of the reference's `do_rollout`
(safe_exploration/episode_runner.py:165-324) without its plotting / sampling-verification branches, and the same loop
for M scenarios in lock-step through ONE batched solve per step (`B200CemSafeMPC.get_action_batch`, BASELINE config
"batched independent MPC solves") -- what `run_episodic`'s per-scenario loop (:37-139) does one scenario at a time.

Environments and the metrics sink are the caller's (the reference's `Environment` and `SacredAggregatedMetrics` fit):
    env.reset(mean, std) -> state;  env.step(action) -> (action, next_state, observation, done, env_result);
    env.n_s, env.n_u;  optional env.collect_metrics();  optional env.random_action()
    metrics.log_scalar(name, value, step);  metrics.log_non_scalars(dict, step)
"""
import time
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np

from .safempc import MpcResult

RolloutResult = Tuple[np.ndarray, np.ndarray, list, np.ndarray, bool]


def _log_episode(metrics, episode_id: int, n_successful: int, mpc_results: List[MpcResult], time_in_solver: float,
                 env_result, env, solver) -> None:
    """episode_runner.py:316-324: the per-episode metrics, under the reference's names."""
    if metrics is None:
        return
    metrics.log_scalar('episode_length', n_successful, episode_id)
    metrics.log_scalar('mpc_found_solution_count', mpc_results.count(MpcResult.FOUND_SOLUTION), episode_id)
    metrics.log_scalar('mpc_previous_solution_count', mpc_results.count(MpcResult.PREVIOUS_SOLUTION), episode_id)
    metrics.log_scalar('safe_controller_fallback_count', mpc_results.count(MpcResult.SAFE_CONTROLLER), episode_id)
    metrics.log_scalar('mean_time_in_solver', float(time_in_solver) / max(1, n_successful), episode_id)
    metrics.log_scalar('env_result', env_result, episode_id)
    if hasattr(env, 'collect_metrics'):
        metrics.log_non_scalars(env.collect_metrics(), episode_id)
    if solver is not None:
        metrics.log_non_scalars(solver.collect_metrics(), episode_id)


def do_rollout(env, n_steps: int, scenario_id: int = 0, episode_id: int = 0, metrics=None, solver=None,
               cost: Optional[Callable] = None, verbosity: int = 0, mean=None, std=None, **_plot_options
               ) -> RolloutResult:
    """One episode (episode_runner.py:165-324).  -> (xx [1 + steps, n_s + n_u], yy [1 + steps, n_s], cc, exit_codes,
    safety_failure): like the reference, row 0 of xx / yy / exit_codes is the zero row it starts its stacks with.
    Plotting / rendering keyword arguments of the reference are accepted and ignored."""
    state = env.reset(mean, std)
    xx = [np.zeros(env.n_s + env.n_u)]
    yy = [np.zeros(env.n_s)]
    exit_codes = [0.0]
    cc, mpc_results = [], []
    total_time_in_solver, n_successful, env_result, safety_failure = 0.0, 0, -1, False
    for _ in range(n_steps):
        if solver is None:
            action, exit_code = env.random_action(), 5
        else:
            t0 = time.time()
            action, mpc_result = solver.get_action(state)
            t_solver = time.time() - t0
            total_time_in_solver += t_solver
            exit_code = 1 if mpc_result in (MpcResult.FOUND_SOLUTION, MpcResult.PREVIOUS_SOLUTION) else 0
            mpc_results.append(mpc_result)
            if verbosity > 0:
                print(f'total time solver in s: {t_solver}')
        action, next_state, observation, done, env_result = env.step(action)
        if cost is not None:
            cc.append(cost(next_state))
        xx.append(np.hstack((state, action)))
        yy.append(np.asarray(observation))
        exit_codes.append(exit_code)
        n_successful += 1
        state = next_state
        if done:
            safety_failure = True
            break
    _log_episode(metrics, episode_id, n_successful, mpc_results, total_time_in_solver, env_result, env, solver)
    return np.vstack(xx), np.vstack(yy), cc, np.asarray(exit_codes).reshape(-1, 1), safety_failure


def do_rollout_batch(envs: Sequence, n_steps: int, solver, scenario_ids: Optional[Sequence[int]] = None,
                     episode_id: int = 0, metrics: Optional[Sequence] = None, cost: Optional[Callable] = None,
                     verbosity: int = 0, mean=None, std=None) -> List[RolloutResult]:
    """M episodes in lock-step: every step is ONE `solver.get_action_batch(states [M, n_s])` -- one device call for all
    scenarios -- followed by the M environment steps.  Scenario i sees exactly the state machine of `get_action`
    (FOUND / PREVIOUS / SAFE_CONTROLLER, safempc_cem.py:231-263) on its own stored plan, so its episode equals a
    `do_rollout` of that scenario alone whose solver draws the same candidates.  A scenario that ends early (`done`)
    keeps its slot in the batch (its last state is re-sent, the result ignored): the batch shape, and with it the
    library's captured CUDA graph, stays the same.

    :param metrics: None, or one metrics sink per scenario (the reference creates one run per scenario)
    :returns: one (xx, yy, cc, exit_codes, safety_failure) per scenario, as `do_rollout` returns them;
              `mean_time_in_solver` is the batched call's time divided by the number of scenarios still running."""
    m = len(envs)
    scenario_ids = list(range(m)) if scenario_ids is None else list(scenario_ids)
    states = [np.asarray(env.reset(mean, std), dtype=np.float64) for env in envs]
    xx = [[np.zeros(env.n_s + env.n_u)] for env in envs]
    yy = [[np.zeros(env.n_s)] for env in envs]
    exit_codes = [[0.0] for _ in envs]
    cc = [[] for _ in envs]
    mpc_results = [[] for _ in envs]
    time_in_solver = [0.0] * m
    n_successful = [0] * m
    env_result = [-1] * m
    failed = [False] * m
    running = [True] * m
    for _ in range(n_steps):
        if not any(running):
            break
        t0 = time.time()
        actions, results = solver.get_action_batch(np.stack(states))
        t_solver = time.time() - t0
        share = t_solver / sum(running)
        if verbosity > 0:
            print(f'total time solver in s: {t_solver} for {sum(running)} scenarios')
        for i, env in enumerate(envs):
            if not running[i]:
                continue
            time_in_solver[i] += share
            mpc_results[i].append(results[i])
            exit_code = 1 if results[i] in (MpcResult.FOUND_SOLUTION, MpcResult.PREVIOUS_SOLUTION) else 0
            action, next_state, observation, done, env_result[i] = env.step(actions[i])
            if cost is not None:
                cc[i].append(cost(next_state))
            xx[i].append(np.hstack((states[i], action)))
            yy[i].append(np.asarray(observation))
            exit_codes[i].append(exit_code)
            n_successful[i] += 1
            states[i] = np.asarray(next_state, dtype=np.float64)
            if done:
                failed[i], running[i] = True, False
    out = []
    for i, env in enumerate(envs):
        _log_episode(None if metrics is None else metrics[i], episode_id, n_successful[i], mpc_results[i],
                     time_in_solver[i], env_result[i], env, solver)
        out.append((np.vstack(xx[i]), np.vstack(yy[i]), cc[i], np.asarray(exit_codes[i]).reshape(-1, 1), failed[i]))
    return out
