#!/usr/bin/env python
"""bench.py -- headline benchmark of the safe CEM-MPC hot path (BASELINE.json metric).

A "step" is one MPC solve (one `ConstrainedCemMpc.get_actions`: I CEM iterations x N candidates x T ellipsoid steps)
on synthetic data of the BASELINE config C2 (InvertedPendulum, GP on 2000 training points, horizon 10, population
16384 per GPU).  metric = candidate ellipsoid-rollout steps / s = N T I / t_solve, whole job.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--dtype f64|f32] [--config C2|C1|C3]

N > 1 is launched by the driver under torch.distributed.run (one rank per GPU): the population is sharded over the
ranks (weak scaling: 16384 candidates per GPU) with one NCCL all-gather of candidate costs per CEM iteration.  The N > 1
line also carries two side records measured in the same process: `strong_scaling` (BASELINE config C3: CartPole, n_train
5000, horizon 15, population 65536 split over the N ranks) and `c4` (BASELINE config C4: independent pendulum MPC
instances x population 4096, instance-sharded, 128 instances per GPU, results gathered on rank 0).  The N = 1 line
carries a `c1` record: BASELINE config C1 (the reference's own CPU-sized case) with the CPU path timed in full beside it.
`--impl reference` times the reference's CPU path (the float64 oracle port, all host threads) on a bounded sample of the
same workload and prints the same JSON line with "impl": "reference".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = 'candidate ellipsoid-rollout steps/sec'
UNIT = 'candidate-steps/s'


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--dtype', default='f64', choices=['f64', 'f32'])
    ap.add_argument('--config', default='C2', choices=['C1', 'C2', 'C3'])
    ap.add_argument('--gemm', default='auto', choices=['auto', 'simt', 'mma', 'tc'])
    ap.add_argument('--tc-digits', type=int, default=0, choices=[0, 3, 4, 5, 6, 7, 8],
                    help='digit planes of the tcgen05 variance GEMM (0 = chosen from outputscale / noise)')
    ap.add_argument('--cpu-sample', type=int, default=2048, help='candidates in the CPU-baseline sample')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-side-records', action='store_true', help='skip the c1 / strong_scaling / c4 records')
    return ap.parse_args()


def make_workload(name: str, world: int):
    from safe_exploration_b200 import workloads
    if name == 'C1':
        return workloads.pendulum_workload(n_train=100, num_rollouts=500 * world if world > 1 else 500, time_horizon=5,
                                           num_elites=20, num_iterations=8, name='C1')
    if name == 'C3':
        return workloads.cartpole_workload(num_rollouts=max(8192 * world, 8192))
    return workloads.pendulum_workload(num_rollouts=16384 * world)


def workload_desc(w, world, dtype):
    return {'workload': f'{w.name}: {w.meta["env"]} safe CEM-MPC, exact GP n_train={w.n_train}, horizon {w.time_horizon}, '
                        f'population {w.num_rollouts} ({w.num_rollouts // world}/GPU), elites {w.num_elites}, '
                        f'iterations {w.num_iterations}',
            'n_s': w.n_s, 'n_u': w.n_u, 'n_train': w.n_train, 'time_horizon': w.time_horizon,
            'population': w.num_rollouts, 'elites': w.num_elites, 'iterations': w.num_iterations,
            'parallelism': f'population-sharded x{world}' if world > 1 else 'single GPU',
            'cache': 'inputs_larger_than_l2 (kernel rows K* of one horizon step: %.0f MB vs 126 MB L2)' %
                     (w.n_s * (w.num_rollouts // world) * w.n_train * (8 if dtype == 'f64' else 4) / 1e6)}


# ---- CPU baseline: the reference's path, float64, eager torch, all host threads ------------------------------------------
def cpu_reference_run(w, sample: int, repeats: int, warmup: int, loops: bool = True):
    """One CEM iteration (sample -> T reachability steps -> constraint loop -> elite selection -> refit) of the oracle
    port with the reference's per-item Python loops, on `sample` candidates of the same workload (full n_train, full
    horizon).  loops=False runs the same math with the three per-item loops (eig, trace, constraints) batched: the
    "reference math, de-looped" line of SURVEY 8(d).  Returns (candidate-steps/s, seconds per iteration, threads)."""
    from oracle import oracle
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    T = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64))
    gp = oracle.OracleGpCemSSM(w.n_s, w.n_u, w.lengthscale, w.outputscale, w.noise)
    gp.update_model(T(w.x_train), T(w.y_train), replace_old=True)
    prob = oracle.SafeCemProblem(gp, w.l_mu, w.l_sigma, w.k_fb, w.a, w.b, w.beta_safety, w.h_mat, w.h_vec, w.act_a,
                                 w.act_b)
    elites = max(2, min(w.num_elites, sample // 4))
    times = []
    for rep in range(warmup + repeats):
        noise = T(oracle.philox_normal_noise(rep, 0, 0, sample, w.time_horizon, w.n_u)).unsqueeze(0)
        t0 = time.perf_counter()
        oracle.cem_solve(prob, T(w.p0), noise, elites, loops=loops)
        dt = time.perf_counter() - t0
        if rep >= warmup:
            times.append(dt)
    sec = float(np.median(times))
    return sample * w.time_horizon / sec, sec, threads


def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    w = make_workload(args.config, 1)
    value, sec, threads = cpu_reference_run(w, args.cpu_sample, max(1, args.steps), max(0, min(args.warmup, 1)))
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': sec * 1e3, 'ms_per_sample': sec * 1e3,
        'step_definition': f'one step of this arm = one bounded SAMPLE of the workload ({args.cpu_sample} candidates x '
                           f'1 CEM iteration at full n_train and horizon), not a whole solve; value = candidate-steps '
                           f'of the sample / its time',
        'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic', 'config': workload_desc(w, 1, 'f64'),
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': threads, 'kind': 'port',
                         'sample': f'{args.cpu_sample} candidates x horizon {w.time_horizon} x 1 CEM iteration at full '
                                   f'n_train={w.n_train} (reference per-item Python loops kept); '
                                   f'the reference itself is Python + gpytorch 0.3.2 and cannot run on this box'},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


# ---- clocks ----------------------------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index: int):
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        if not self._run_nvml():
            self._run_smi()

    def _run_nvml(self) -> bool:
        """In-process NVML (nvidia_ml_py): ~1 ms per sample, so even a 0.5 s timed region is sampled dozens of times."""
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get('CUDA_VISIBLE_DEVICES')
            idx = self.index
            if vis and len(vis.split(',')) > self.index and vis.split(',')[self.index].strip().isdigit():
                idx = int(vis.split(',')[self.index])
            h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            const = lambda name: getattr(nv, 'nvmlClocksEventReason' + name, None) or \
                getattr(nv, 'nvmlClocksThrottleReason' + name)
            bits = {'hw_slowdown': const('HwSlowdown'), 'hw_thermal_slowdown': const('HwThermalSlowdown'),
                    'sw_thermal_slowdown': const('SwThermalSlowdown'), 'sw_power_cap': const('SwPowerCap')}
            get_reasons = getattr(nv, 'nvmlDeviceGetCurrentClocksEventReasons', None) or \
                nv.nvmlDeviceGetCurrentClocksThrottleReasons
        except Exception:
            return False
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                mask = int(get_reasons(h))
                for n, bit in bits.items():
                    if mask & int(bit):
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop.wait(0.02)
        return True

    def _run_smi(self):
        q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        while not self._stop.is_set():
            try:
                out = subprocess.run(['nvidia-smi', f'--id={self.index}', f'--query-gpu={q}',
                                      '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
                f = [s.strip() for s in out.strip().split(',')]
                self.samples.append(float(f[0]))
                self.max_mhz = float(f[1])
                for n, v in zip(names, f[2:6]):
                    if v.lower().startswith('active'):
                        self.reasons.add(n)
            except Exception:
                pass
            self._stop.wait(0.2)

    def __enter__(self):
        self._thread.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._thread.join(timeout=6)

    def summary(self):
        return {'sm_mhz': float(np.median(self.samples)) if self.samples else None, 'sm_max_mhz': self.max_mhz,
                'reasons': sorted(self.reasons), 'samples': len(self.samples)}


def profiled_traffic(kernel_name: str, workload: str):
    """DRAM bytes (read + write) per launch of the dominant kernel from the committed `ncu --set full` capture
    (profiles/ncu_traffic.json, written by tools/ncu_traffic.py from the capture's raw page together with the commit,
    kernel name and workload it was taken on).  Returns (bytes | None, stamp): None when no capture matches the kernel
    that ran -- a stale figure is worse than none."""
    path = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')
    try:
        with open(path) as f:
            entries = json.load(f)
    except Exception:
        return None, None
    for e in entries:
        if e.get('kernel') == kernel_name and e.get('workload') == workload:
            return e.get('dram_bytes_per_launch'), {k: e.get(k) for k in ('commit', 'kernel', 'workload', 'source')}
    return None, {'note': f'no committed capture of {kernel_name} on {workload}'}


def measured_peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), 'measured'
    return {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0}, 'fallback'


# ---- the B200 arm ----------------------------------------------------------------------------------------------------------
def _timed_solves(lib, _lib, ssm, mpc, p0, steps, barrier, dev, instrument, seed=None):
    """K solves between two CUDA events on the solve stream.  instrument=False: the production path (CUDA graphs, fused
    steps); instrument=True: the library brackets every variance-GEMM / kernel-row launch with events
    (se_gp_timing_enable), which switches graphs and step fusion off -- used only for the per-kernel split."""
    import ctypes as C
    if instrument:
        _lib.check(lib.se_gp_timing_enable(ssm.handle, 1))
    lib.se_launch_count_reset()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    res = None
    for _ in range(steps):
        res = mpc.solve(p0, sync=False, seed=seed)
    ev1.record()
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    out = dict(ms_total=ms_total, launches=int(lib.se_launch_count()), res=res)
    if instrument:
        gemm_ms, kstar_ms = C.c_double(0), C.c_double(0)
        gemm_n, kstar_n = C.c_int64(0), C.c_int64(0)
        _lib.check(lib.se_gp_timing_read(ssm.handle, C.byref(gemm_ms), C.byref(gemm_n), C.byref(kstar_ms),
                                         C.byref(kstar_n), _lib.stream_ptr(dev)))
        _lib.check(lib.se_gp_timing_enable(ssm.handle, 0))
        out.update(gemm_ms=gemm_ms.value, gemm_n=int(gemm_n.value), kstar_ms=kstar_ms.value, kstar_n=int(kstar_n.value))
    return out


def _max_over_ranks(x, world, dev):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _gemm_kernel_name(digits, rows_local=0):
    """Name of the variance-GEMM kernel the library launches for this run (csrc/inst_ozaki.cu: launch_trisumsq_ozaki)."""
    mode = int(os.environ.get('SE_OZ_REUSE', '2'))
    shape = int(os.environ.get('SE_OZ_CLUSTER', '12'))
    tiles = (rows_local + 127) // 128
    if mode == 2 and shape in (12, 21) and not (shape == 12 and tiles % 2):
        return f'oz::gp_trisumsq_ozaki_cluster_kernel<{digits},{shape // 10},{shape % 10}>'
    return f'oz::gp_trisumsq_ozaki_kernel<{digits},{mode}>'


def side_record_c1(lib, _lib, workloads, dev, dtype, barrier, with_cpu):
    """BASELINE config C1 (the only config the reference runs on its own CPU path today): GPU solve time, and the CPU
    path (oracle port with the reference's per-item loops, all host threads) timed IN FULL, median of 5 solves
    (BASELINE.md section 2)."""
    w = make_workload('C1', 1)
    ssm, mpc = workloads.build_solver(w, dtype=dtype, device=dev, record='none', seed=77)
    p0 = torch.as_tensor(w.p0, device=dev, dtype=dtype).unsqueeze(0)
    for _ in range(5):
        mpc.solve(p0, sync=False)
    t = _timed_solves(lib, _lib, ssm, mpc, p0, 50, barrier, dev, False)
    ms = t['ms_total'] / 50
    cand_steps = w.num_rollouts * w.time_horizon * w.num_iterations
    rec = {'workload': workload_desc(w, 1, 'f64')['workload'], 'ms_per_solve': ms, 'value': cand_steps / (ms * 1e-3),
           'unit': UNIT, 'solves_timed': 50, 'gpu_launches_per_solve': t['launches'] / 50}
    if with_cpu:
        from oracle import oracle
        threads = os.cpu_count() or 1
        torch.set_num_threads(threads)
        T = lambda x: torch.as_tensor(np.asarray(x, dtype=np.float64))
        gp = oracle.OracleGpCemSSM(w.n_s, w.n_u, w.lengthscale, w.outputscale, w.noise)
        gp.update_model(T(w.x_train), T(w.y_train), replace_old=True)
        prob = oracle.SafeCemProblem(gp, w.l_mu, w.l_sigma, w.k_fb, w.a, w.b, w.beta_safety, w.h_mat, w.h_vec, w.act_a,
                                     w.act_b)
        times = []
        for rep in range(6):
            noise = T(np.stack([oracle.philox_normal_noise(rep, i, 0, w.num_rollouts, w.time_horizon, w.n_u)
                                for i in range(w.num_iterations)]))
            t0 = time.perf_counter()
            oracle.cem_solve(prob, T(w.p0), noise, w.num_elites, loops=True)
            if rep > 0:
                times.append(time.perf_counter() - t0)
        sec = float(np.median(times))
        rec['cpu_baseline'] = {'value': cand_steps / sec, 'unit': UNIT, 'cores': threads, 'kind': 'port',
                               'ms_per_solve': sec * 1e3,
                               'sample': 'the whole C1 solve (500 candidates x horizon 5 x 8 iterations), median of 5, '
                                         'float64 eager torch, reference per-item loops kept'}
        rec['gpu_over_cpu'] = sec * 1e3 / ms
    mpc._destroy()
    return rec


def side_record_strong_scaling(lib, _lib, workloads, sedist, dev, dtype, rank, world, barrier):
    """BASELINE config C3 at its full size with the population split over the ranks (strong scaling): CartPole, n_train
    5000, horizon 15, population 65536, 8 iterations; 3 timed solves + one instrumented solve for the kernel split."""
    w = workloads.cartpole_workload(num_rollouts=65536)
    ssm, mpc = workloads.build_solver(w, dtype=dtype, device=dev, record='none', seed=4321)
    comm = sedist.attach_population_sharding(mpc) if world > 1 else 0
    p0 = torch.as_tensor(w.p0, device=dev, dtype=dtype).unsqueeze(0)
    mpc.solve(p0, sync=False)
    t = _timed_solves(lib, _lib, ssm, mpc, p0, 3, barrier, dev, False)
    ms = _max_over_ranks(t['ms_total'], world, dev) / 3
    ti = _timed_solves(lib, _lib, ssm, mpc, p0, 1, barrier, dev, True)
    cand_steps = w.num_rollouts * w.time_horizon * w.num_iterations
    impl, digits = ssm.gemm_info()
    rec = {'workload': workload_desc(w, world, 'f64')['workload'], 'scaling': 'strong', 'n_gpus': world,
           'ms_per_solve': ms, 'value': cand_steps / (ms * 1e-3), 'unit': UNIT, 'solves_timed': 3,
           'gemm_impl': impl, 'variance_digit_planes': digits,
           'gemm_share_of_solve': ti['gemm_ms'] / ti['ms_total'], 'kstar_share_of_solve': ti['kstar_ms'] / ti['ms_total'],
           'gemm_avg_launch_ms': ti['gemm_ms'] / max(1, ti['gemm_n']),
           'found_solution': bool(int(t['res']['found'][0]))}
    barrier()
    mpc._destroy()
    if world > 1:
        sedist.destroy_nccl_comm(comm)
    ssm._destroy()
    return rec


def side_record_c4(lib, _lib, workloads, sedist, dev, dtype, rank, world, barrier):
    """BASELINE config C4: independent pendulum MPC instances x population 4096, instance-sharded (no communication on
    the data path), 128 instances per GPU -- 1024 instances on 8 GPUs --, results gathered on every rank at the end."""
    per_gpu = 128
    M = per_gpu * world
    w = workloads.pendulum_workload(num_rollouts=4096)
    ssm, mpc = workloads.build_solver(w, dtype=dtype, device=dev, record='none', seed=99, num_instances=per_gpu)
    p0_all = 0.05 * torch.randn((M, w.n_s), generator=torch.Generator().manual_seed(3), dtype=torch.float64)
    sedist.solve_instances_sharded(mpc, p0_all, seed=10)  # warm-up (workspaces, digit image)
    barrier()
    t0 = time.perf_counter()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    plans, found = sedist.solve_instances_sharded(mpc, p0_all, seed=11)
    ev1.record()
    barrier()
    wall = time.perf_counter() - t0
    ms = _max_over_ranks(ev0.elapsed_time(ev1), world, dev)
    cand_steps = M * w.num_rollouts * w.time_horizon * w.num_iterations
    rec = {'workload': f'C4: {M} independent InvertedPendulum MPC instances ({per_gpu}/GPU) x population 4096, exact GP '
                       f'n_train={w.n_train}, horizon {w.time_horizon}, elites {w.num_elites}, iterations '
                       f'{w.num_iterations}', 'scaling': 'weak (instance-sharded, no data-path collective)',
           'n_gpus': world, 'instances': M, 'ms_per_batch_solve': ms, 'ms_per_instance': ms / per_gpu,
           'value': cand_steps / (ms * 1e-3), 'unit': UNIT, 'wall_s_including_gather': wall,
           'instances_with_a_feasible_plan': int(found.sum()), 'gathered_plans_shape': list(plans.shape)}
    mpc._destroy()
    ssm._destroy()
    return rec


def run_b200(args):
    import ctypes as C
    import torch.distributed as dist
    import safe_exploration_b200 as se
    from safe_exploration_b200 import _lib, workloads
    from safe_exploration_b200 import dist as sedist

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if world != args.gpus and world > 1:
        raise SystemExit(f'--gpus {args.gpus} but WORLD_SIZE={world}')
    torch.cuda.set_device(local_rank)
    dev = f'cuda:{local_rank}'
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device(dev))
    lib = se.load_library()
    dtype = torch.float64 if args.dtype == 'f64' else torch.float32
    gemm = {'auto': _lib.SE_GEMM_AUTO, 'simt': _lib.SE_GEMM_SIMT, 'mma': _lib.SE_GEMM_MMA, 'tc': _lib.SE_GEMM_TC}[args.gemm]

    w = make_workload(args.config, world)
    ssm, mpc = workloads.build_solver(w, dtype=dtype, device=dev, record='none', seed=1234, gemm_impl=gemm)
    if args.tc_digits:
        ssm.set_tc_digits(args.tc_digits)
    comm = sedist.attach_population_sharding(mpc) if world > 1 else 0
    p0 = torch.as_tensor(w.p0, device=dev, dtype=dtype).unsqueeze(0)
    cand_steps = w.num_rollouts * w.time_horizon * w.num_iterations  # whole job, all ranks

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # warm-up (also grows every workspace, so the timed region makes no allocation inside the library)
    warm = max(3, args.warmup)
    for _ in range(warm):
        mpc.solve(p0, sync=False)
    barrier()

    # ---- device-resident timing of the PRODUCTION path: K solves between two events, max over ranks ----
    with ClockSampler(local_rank) as clocks:
        plain = _timed_solves(lib, _lib, ssm, mpc, p0, args.steps, barrier, dev, False)
    ms_total = _max_over_ranks(plain['ms_total'], world, dev)
    ms_per_step = ms_total / args.steps
    value = cand_steps / (ms_per_step * 1e-3)
    launches = plain['launches']
    res = plain['res']
    flags = res['flags'].cpu().tolist()

    # ---- second pass, instrumented: per-kernel split (models of <= 128 training points run one fused rollout kernel
    # per CEM iteration, which has no separate variance GEMM to time) ----
    fused_small = w.n_train <= 128 and args.gemm == 'auto' and os.environ.get('SE_ROLLOUT_FUSED', '1') != '0'
    inst = None
    if not fused_small:
        inst = _timed_solves(lib, _lib, ssm, mpc, p0, args.steps, barrier, dev, True)

    # ---- end to end through the public API: host state in, host action out, every step ----
    class Env:
        n_s, n_u = w.n_s, w.n_u
        l_mu, l_sigm = w.l_mu, w.l_sigma
        objective_spec = None

    class Conf:
        mpc_time_horizon, cem_num_rollouts = w.time_horizon, w.num_rollouts
        cem_num_elites, cem_num_iterations = w.num_elites, w.num_iterations
        use_prior_model, cem_record = True, 'none'

    solver = se.B200CemSafeMPC(ssm, mpc._constraints, Env, Conf, {'lin_model': (w.a, w.b)}, None, None, w.beta_safety,
                               safe_policy=lambda s: np.zeros(w.n_u), lqr=_FixedGain(w.k_fb), mpc=mpc, verbose=False)
    state_host = np.ascontiguousarray(w.p0, dtype=np.float64)
    for _ in range(2):
        solver.get_action(state_host)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        action, result = solver.get_action(state_host)
    barrier()
    e2e_sec = _max_over_ranks((time.perf_counter() - t0) / args.steps, world, dev)
    esz = 8 if args.dtype == 'f64' else 4
    e2e = {'value': cand_steps / e2e_sec, 'unit': UNIT, 'ms_per_solve': e2e_sec * 1e3,
           'h2d_bytes_per_step': w.n_s * esz, 'd2h_bytes_per_step': w.time_horizon * w.n_u * esz + 8 + 16,
           'api': 'B200CemSafeMPC.get_action(ndarray) -> (ndarray, MpcResult)'}

    # ---- roofline of the dominant kernel (variance GEMM) ----
    peaks, peak_src = measured_peaks()
    impl_name, digits = ssm.gemm_info()
    rows_local = w.num_rollouts // world
    fp64_flop_per_launch = w.n_s * float(rows_local) * float(w.n_train) ** 2  # n^2/2 FMA per (candidate, output)

    def pipe_peak(kinds):
        best, f = 0.0, C.c_double(0)
        for k in kinds:
            _lib.check(lib.se_measure_peak(k, C.byref(f)))
            best = max(best, f.value)
        return best / 1e12

    fp64_peak = pipe_peak([1, 2]) if (rank == 0 and args.dtype == 'f64') else None
    roofline = None
    if inst is not None:
        gemm_avg_ms = inst['gemm_ms'] / max(1, inst['gemm_n'])
        extra = {}
        if impl_name == 'tcgen05-i8':
            # every float64 FMA of the contraction is evaluated as S (S + 1) / 2 exact int8 digit-plane MACs
            pairs = digits * (digits + 1) // 2
            work_per_launch = pairs * fp64_flop_per_launch  # int8 OP (2 x MAC)
            peak_burst = pipe_peak([3]) if rank == 0 else None
            peak = pipe_peak([4]) if rank == 0 else None  # the GEMM runs inside a long step: sustained figure
            pipe = f'int8 tensor (tcgen05.mma kind::i8, {digits} radix-256 digit planes, {pairs} MMAs per K step)'
            unit = 'TOP/s'
            peak_src_txt = ('int8 tcgen05 peak measured in this run with se_measure_peak(4): back-to-back 128x256x32 MMAs, '
                            'operands resident, sustained for 1.5 s under the power cap; the other denominators are in '
                            '`frac_against`')
            achieved = work_per_launch / (gemm_avg_ms * 1e-3) / 1e12 if gemm_avg_ms > 0 else None
            if achieved and rank == 0:
                extra['frac_against'] = {
                    'int8_sustained_measured_in_run': achieved / peak if peak else None,
                    'int8_burst_measured_in_run': achieved / peak_burst if peak_burst else None,
                    f'2x_bf16_sustained_MEASURED_PEAKS_{peak_src}': achieved / (2 * peaks['bf16_tflops_sustained'])
                    if peaks.get('bf16_tflops_sustained') else None,
                    f'2x_bf16_burst_MEASURED_PEAKS_{peak_src}': achieved / (2 * peaks['bf16_tflops'])
                    if peaks.get('bf16_tflops') else None,
                    'int8_nominal_4500': achieved / 4500.0}
                extra['peaks_top_s'] = {'int8_sustained_in_run': peak, 'int8_burst_in_run': peak_burst,
                                        '2x_bf16_sustained': 2 * peaks.get('bf16_tflops_sustained', 0),
                                        '2x_bf16_burst': 2 * peaks.get('bf16_tflops', 0), 'int8_nominal': 4500.0}
        else:
            work_per_launch = fp64_flop_per_launch
            peak = (fp64_peak if args.dtype == 'f64' else pipe_peak([0])) if rank == 0 else None
            pipe = 'FP64 (DMMA / DFMA)' if args.dtype == 'f64' else 'FP32 FFMA'
            unit = 'TFLOP/s'
            peak_src_txt = ('measured in this run with se_measure_peak (register-resident FMA / DMMA loop on every SM); '
                            f'MEASURED_PEAKS.json ({peak_src}) holds only HBM and bf16 numbers')
            achieved = work_per_launch / (gemm_avg_ms * 1e-3) / 1e12 if gemm_avg_ms > 0 else None
        fp64_equiv = fp64_flop_per_launch / (gemm_avg_ms * 1e-3) / 1e12 if gemm_avg_ms > 0 else None
        op_bytes = digits if impl_name == 'tcgen05-i8' else esz
        traffic, stamp = (profiled_traffic(_gemm_kernel_name(digits, rows_local), f'{args.config}/N{rows_local}')
                          if impl_name == 'tcgen05-i8' else (None, None))
        roofline = {
            'bound': 'tensor', 'kernel': 'gp_trisumsq (variance GEMM: || L^-1 k* ||^2)',
            'kernel_name': _gemm_kernel_name(digits, rows_local) if impl_name == 'tcgen05-i8' else impl_name, 'impl': impl_name,
            'pipe': pipe, 'achieved': achieved, 'peak': peak, 'unit': unit,
            'frac': (achieved / peak) if (achieved and peak) else None, 'peak_source': peak_src_txt,
            'work_per_launch': work_per_launch, 'avg_launch_ms': gemm_avg_ms, 'launches': inst['gemm_n'],
            'timed_in': 'second, instrumented pass (events around every launch; graphs / step fusion off): '
                        f'{inst["ms_total"] / args.steps:.3f} ms per solve vs {plain["ms_total"] / args.steps:.3f} ms '
                        'on the production path that `value` is taken from',
            'fp64_equivalent_tflops': fp64_equiv, 'fp64_pipe_peak_tflops': fp64_peak,
            'share_of_step': inst['gemm_ms'] / inst['ms_total'] if inst['ms_total'] > 0 else None,
            'kstar_share_of_step': inst['kstar_ms'] / inst['ms_total'] if inst['ms_total'] > 0 else None,
            'kstar_avg_launch_ms': inst['kstar_ms'] / max(1, inst['kstar_n']),
            'traffic': traffic, 'traffic_source': stamp,
            'hbm': {'algorithmic_bytes_per_launch': (w.n_s * rows_local * w.n_train + w.n_s * w.n_train ** 2 / 2) * op_bytes,
                    'peak_gbs': peaks.get('hbm_gbs')},
        }
        roofline.update(extra)
    else:
        # no separate variance GEMM: the whole rollout is one latency-bound kernel; quote the solve's FP64 work rate
        f_step = w.n_s * (w.n_train ** 2 + w.n_train * (5 * (w.n_s + w.n_u) + 8))
        ach = (cand_steps / world) * f_step / (ms_per_step * 1e-3) / 1e12
        roofline = {'bound': 'tensor', 'kernel': 'rollout_small_kernel (kernel rows + variance + ellipsoid step of a whole '
                    'rollout, one CTA per candidate)', 'impl': 'fused-rollout', 'pipe': 'FP64 (DFMA)', 'achieved': ach,
                    'peak': fp64_peak, 'unit': 'TFLOP/s', 'frac': (ach / fp64_peak) if fp64_peak else None,
                    'peak_source': 'se_measure_peak (register-resident DFMA / DMMA loop on every SM)',
                    'note': 'latency-bound by construction: N x T x I = %d candidate-steps of %.2e flop each; the figure '
                            'is algorithmic flop of the whole solve over its time' % (cand_steps, f_step),
                    'work_per_launch': (cand_steps / world / w.num_iterations) * f_step, 'traffic': None}

    # ---- accuracy of the variance path actually timed, against the float64 DMMA kernel, on 2048 visited inputs ----
    variance_check = None
    if rank == 0 and args.dtype == 'f64' and impl_name != 'dmma':
        gen = torch.Generator(device='cpu').manual_seed(7)
        zs = torch.randn(2048, w.n_s + w.n_u, generator=gen, dtype=torch.float64) * 0.15
        zs = zs.to(dev)
        _, v_run = ssm.predict_without_jacobians(zs[:, :w.n_s], zs[:, w.n_s:])
        ssm.set_gemm_impl(_lib.SE_GEMM_MMA)
        _, v_ref = ssm.predict_without_jacobians(zs[:, :w.n_s], zs[:, w.n_s:])
        ssm.set_gemm_impl(gemm)
        rel = ((v_run - v_ref).abs() / v_ref.abs())
        variance_check = {'against': 'float64 DMMA kernel', 'inputs': 2048, 'max_rel_err': float(rel.max()),
                          'median_rel_err': float(rel.median())}

    accuracy = None
    if impl_name == 'tcgen05-i8':
        accuracy = {'variance_digit_planes': digits,
                    'accuracy_class': 'float64 (variance to 1e-8 relative)' if digits >= 6 else
                    'reduced: %d digit planes (float32-tolerance dial of se_gp_set_tc_digits; achieved error in '
                    'variance_check)' % digits}
    cpu_baseline = None
    if rank == 0 and not args.no_cpu_baseline:
        w1 = make_workload(args.config, 1)
        v, sec, threads = cpu_reference_run(w1, args.cpu_sample, 2, 1)
        cpu_baseline = {'value': v, 'unit': UNIT, 'cores': threads, 'kind': 'port',
                        'sample': f'{args.cpu_sample} candidates x horizon {w1.time_horizon} x 1 CEM iteration at full '
                                  f'n_train={w1.n_train}, float64 eager torch, reference per-item loops kept '
                                  f'({sec:.2f} s per sample)'}
        # the same math with the reference's per-item Python loops batched, so the ratio is not credited to loop removal
        v2, sec2, _ = cpu_reference_run(w1, args.cpu_sample, 2, 1, loops=False)
        cpu_baseline['delooped'] = {'value': v2, 'unit': UNIT, 'cores': threads, 'kind': 'port',
                                    'sample': f'same sample, eig / trace / constraint loops batched ({sec2:.2f} s)'}

    # the headline solver is done: release it before the side records build their own models
    found_solution, last_result = bool(int(res['found'][0])), result.name
    if world > 1:
        barrier()
    mpc._destroy()
    if world > 1:
        sedist.destroy_nccl_comm(comm)
    del solver

    side = {}
    if not args.no_side_records and args.config == 'C2' and args.dtype == 'f64' and args.gemm == 'auto':
        try:
            if world == 1:
                side['c1'] = side_record_c1(lib, _lib, workloads, dev, dtype, barrier, not args.no_cpu_baseline)
            else:
                side['strong_scaling'] = side_record_strong_scaling(lib, _lib, workloads, sedist, dev, dtype, rank, world,
                                                                    barrier)
                side['c4'] = side_record_c4(lib, _lib, workloads, sedist, dev, dtype, rank, world, barrier)
        except Exception as exc:  # a side record must never cost the headline line
            side['side_record_error'] = f'{type(exc).__name__}: {exc}'

    if rank == 0:
        line = {
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': warm, 'ms_per_step': ms_per_step, 'mpc_solve_ms': ms_per_step,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': args.dtype, 'data': 'synthetic',
            'config': workload_desc(w, world, args.dtype), 'e2e': e2e, 'gpu_launches': launches, 'roofline': roofline,
            'accuracy': accuracy, 'variance_check': variance_check, 'cpu_baseline': cpu_baseline,
            'clocks': clocks.summary(), 'numeric_flags': flags, 'found_solution': found_solution,
            'last_result': last_result,
        }
        line.update(side)
        print(json.dumps(line), flush=True)
    if world > 1:
        barrier()
        dist.destroy_process_group()


class _FixedGain:
    """LqrFeedbackController stand-in carrying the workload's precomputed LQR gain."""

    def __init__(self, k_fb):
        self._k = np.asarray(k_fb, dtype=np.float64)

    def get_control_matrix(self):
        return self._k

    def get_control_matrix_pytorch(self):
        return torch.as_tensor(self._k)


if __name__ == '__main__':
    a = parse_args()
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_b200(a)
