"""Multi-GPU plumbing: one process per GPU, `torch.distributed` for the rendezvous, the data path's one
collective (all-gather of per-candidate costs) issued by libsecem.so itself on the solve stream through NCCL.

Two ways to use W GPUs:
  * population sharding  -- rank r owns candidates [r N/W, (r+1) N/W) of every MPC instance; one all-gather of
    (objective, constraint) per CEM iteration; every rank takes the same elite decision and regenerates the elites'
    action sequences from the counter-based RNG, so nothing else is exchanged.  `attach_population_sharding`.
  * instance sharding    -- independent MPC problems are dealt out to the ranks, no communication on the data path.
    `shard_instances`.
The GP training set and its factors are replicated on every rank.
"""
import ctypes as C
from typing import Optional, Tuple

import torch
import torch.distributed as dist

from . import _lib


def shard_range(total: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous even split used for the population: (first, count).  `total` must be divisible by `world_size`."""
    if total % world_size != 0:
        raise ValueError(f'{total} candidates cannot be split evenly over {world_size} ranks')
    count = total // world_size
    return rank * count, count


def shard_instances(num_instances: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Instance-parallel split (first, count); the remainder goes to the lowest ranks."""
    base, rem = divmod(num_instances, world_size)
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return first, count


def owner_of(global_id: int, total: int, world_size: int) -> Tuple[int, int]:
    """(rank, local index) of a global candidate id under `shard_range`."""
    count = total // world_size
    return global_id // count, global_id % count


def create_nccl_comm(group: Optional[dist.ProcessGroup] = None, device: Optional[torch.device] = None) -> int:
    """Create an ncclComm_t for libsecem.so over the ranks of `group` (default: the world).  Rank 0 makes the unique
    id, torch.distributed carries its 128 bytes to the other ranks (any backend), every rank calls ncclCommInitRank."""
    lib = _lib.load()
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    buf = C.create_string_buffer(128)
    if rank == 0:
        _lib.check(lib.se_nccl_unique_id(buf))
    backend = dist.get_backend(group)
    t = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
    if backend == 'nccl':
        dev = device if device is not None else torch.device('cuda', torch.cuda.current_device())
        t = t.to(dev)
    dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    uid = bytes(t.cpu().numpy().tobytes())
    comm = C.c_void_p(None)
    _lib.check(lib.se_nccl_comm_init(C.byref(comm), world, uid, rank))
    return comm.value


def destroy_nccl_comm(comm: Optional[int]) -> None:
    if comm:
        _lib.load().se_nccl_comm_destroy(C.c_void_p(comm))


def attach_population_sharding(mpc, group: Optional[dist.ProcessGroup] = None) -> int:
    """Shard `mpc`'s population over the ranks of `group`; returns the communicator (keep it alive, destroy it with
    `destroy_nccl_comm`)."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    if world == 1:
        mpc.set_comm(None, 0, 1)
        return 0
    comm = create_nccl_comm(group, mpc._ssm.device)
    mpc.set_comm(comm, rank, world)
    return comm


def solve_instances_sharded(mpc, p0_all: torch.Tensor, seed: Optional[int] = None,
                            group: Optional[dist.ProcessGroup] = None):
    """BASELINE config C4 -- M independent MPC problems dealt out to the ranks of `group` (`shard_instances`), no
    communication on the data path, results gathered on every rank at the end: what the per-scenario loop of
    episode_runner.py:206-231 does one solve at a time.

    :param mpc: this rank's fused B200ConstrainedCemMpc; its `first_instance` / `num_instances` are set here
    :param p0_all: [M, n_s] initial states of ALL instances (host or device; every rank passes the same tensor)
    :returns: plans [M, T, n_u], found [M] (bool) for all instances, in global instance order.  Identical to one batched
              solve of all M instances on one GPU with the same seed (the candidate RNG is keyed by global instance id)."""
    rank, world = (dist.get_rank(group), dist.get_world_size(group)) if dist.is_initialized() else (0, 1)
    M = p0_all.shape[0]
    first, count = shard_instances(M, rank, world)
    dev = mpc._ssm.device
    T_len, n_u = mpc._time_horizon, mpc._action_dimen
    base, rem = divmod(M, world)
    cap = base + (1 if rem else 0)  # every rank contributes a slab of the same size to the gather
    plans = torch.zeros((cap, T_len, n_u), dtype=mpc._ssm.dtype, device=dev)
    found = torch.zeros((cap,), dtype=torch.int32, device=dev)
    if count > 0:
        mpc.first_instance = first
        res = mpc.solve(p0_all[first:first + count].to(device=dev, dtype=mpc._ssm.dtype), seed=seed)
        plans[:count], found[:count] = res['actions'], res['found']
    if world == 1:
        return plans[:M], found[:M].bool()
    backend = dist.get_backend(group)
    send_p, send_f = (plans, found) if backend == 'nccl' else (plans.cpu(), found.cpu())
    all_p = [torch.empty_like(send_p) for _ in range(world)]
    all_f = [torch.empty_like(send_f) for _ in range(world)]
    dist.all_gather(all_p, send_p, group=group)
    dist.all_gather(all_f, send_f, group=group)
    out_p, out_f = [], []
    for r in range(world):
        _, c = shard_instances(M, r, world)
        out_p.append(all_p[r][:c])
        out_f.append(all_f[r][:c])
    return torch.cat(out_p), torch.cat(out_f).bool()


def pack_costs(objective: torch.Tensor, constraint: torch.Tensor) -> torch.Tensor:
    """Local costs [M, N_local] x2 -> the send buffer layout of the device path: [2, M, N_local]
    (plane 0 = objective, plane 1 = constraint; csrc/cem_kernels.cuh `cem_pack_costs_kernel`)."""
    return torch.stack((objective, constraint), dim=0).contiguous()


def unpack_costs(pack_all: torch.Tensor, world_size: int, num_instances: int, n_local: int):
    """Gathered buffer [W, 2, M, N_local] -> (objective [M, N], constraint [M, N]) in GLOBAL candidate order
    (global id g lives on rank g // N_local at local index g % N_local; `CostView::get` in csrc/cem_kernels.cuh)."""
    v = pack_all.reshape(world_size, 2, num_instances, n_local)
    obj = v[:, 0].permute(1, 0, 2).reshape(num_instances, world_size * n_local)
    con = v[:, 1].permute(1, 0, 2).reshape(num_instances, world_size * n_local)
    return obj, con
