"""Multi-GPU layout of the step path: worlds are independent, so rank r owns a contiguous world range and
its own Model/Data; there is no per-step collective.  The only exchange is an optional
all-gather of an observation tensor (qpos | qvel) at rollout boundaries."""

from __future__ import annotations

from typing import Tuple

import torch


def world_shard(nworld_total: int, rank: int, world_size: int) -> Tuple[int, int]:
  """Contiguous world range [w0, w1) of `rank`; the remainder goes to the lowest ranks."""
  if world_size < 1 or not (0 <= rank < world_size):
    raise ValueError(f"invalid rank {rank} for world_size {world_size}")
  if nworld_total < 0:
    raise ValueError("nworld_total must be >= 0")
  base, rem = divmod(nworld_total, world_size)
  w0 = rank * base + min(rank, rem)
  return w0, w0 + base + (1 if rank < rem else 0)


def shard_pool_size(naconmax_total: int, nworld_total: int, w0: int, w1: int) -> int:
  """Size of a rank's contact pool: the global pool split in proportion to its worlds (rounded up)."""
  if nworld_total == 0:
    return 0
  return -(-naconmax_total * (w1 - w0) // nworld_total)


def observation(d) -> torch.Tensor:
  """(nworld, nq + nv) observation of a Data shard."""
  return torch.cat([d.qpos, d.qvel], dim=1)


def all_gather_observation(obs: torch.Tensor, nworld_total: int, group=None) -> torch.Tensor:
  """Gathers per-rank observation shards into the (nworld_total, nobs) tensor, in world order.

  Shards may be ragged by one world (see world_shard), so they are padded to the largest shard for the
  collective and trimmed afterwards.  NCCL (NVLink/NVSwitch) on GPU tensors, gloo on CPU tensors.
  """
  import torch.distributed as dist

  world = dist.get_world_size(group)
  sizes = [world_shard(nworld_total, r, world) for r in range(world)]
  nmax = max(b - a for a, b in sizes)
  pad = torch.zeros((nmax, obs.shape[1]), dtype=obs.dtype, device=obs.device)
  pad[: obs.shape[0]] = obs
  out = torch.empty((world * nmax, obs.shape[1]), dtype=obs.dtype, device=obs.device)
  dist.all_gather_into_tensor(out, pad, group=group)
  return torch.cat([out[r * nmax : r * nmax + (b - a)] for r, (a, b) in enumerate(sizes)], dim=0)
