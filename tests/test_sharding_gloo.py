"""N>1 path on CPU: two gloo ranks each own a world shard, advance it independently (CPU oracle as the
stand-in compute: the CUDA step cannot run here) and all-gather the observation; the result must equal
the unsharded run bit for bit (worlds are independent)."""

import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from mujoco_warp_b200 import sharding
from tests import helpers as H

NWORLD, NSTEP = 7, 3  # ragged on purpose


def test_world_shard_partitions_all_worlds():
  for n in (0, 1, 7, 8, 8192):
    for ws in (1, 2, 3, 8):
      ranges = [sharding.world_shard(n, r, ws) for r in range(ws)]
      assert ranges[0][0] == 0 and ranges[-1][1] == n
      for (a0, a1), (b0, b1) in zip(ranges, ranges[1:]):
        assert a1 == b0 and a1 >= a0
      sizes = [b - a for a, b in ranges]
      assert max(sizes) - min(sizes) <= 1
  with pytest.raises(ValueError):
    sharding.world_shard(8, 2, 2)
  assert sharding.shard_pool_size(196608, 8192, 0, 4096) == 98304
  assert sharding.shard_pool_size(10, 3, 0, 2) == 7


def _free_port():
  s = socket.socket()
  s.bind(("127.0.0.1", 0))
  p = s.getsockname()[1]
  s.close()
  return p


def _rollout(mjm, qpos, qvel, ctrl):
  from oracle import oracle as O

  om, d = H.oracle_data(mjm, qpos, qvel, ctrl, 24, 64)
  for _ in range(NSTEP):
    O.run(om, d, "step")
  return np.concatenate([d.qpos, d.qvel], axis=1)


def _worker(rank, world, port, ret):
  import torch.distributed as dist

  import mujoco_warp_b200 as mjw

  os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
  dist.init_process_group("gloo", rank=rank, world_size=world)
  try:
    mjm = mjw.load_mjcf(H.HUMANOID)
    qpos, qvel, ctrl = H.make_states(mjm, NWORLD, seed=99)
    w0, w1 = sharding.world_shard(NWORLD, rank, world)
    obs = torch.from_numpy(_rollout(mjm, qpos[w0:w1], qvel[w0:w1], ctrl[w0:w1]))
    full = sharding.all_gather_observation(obs, NWORLD)
    # barrier + max-over-ranks reduction as bench.py does for its timing
    t = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.barrier()
    if rank == 0:
      ret["full"] = full.numpy()
      ret["tmax"] = float(t)
  finally:
    dist.destroy_process_group()


def test_two_rank_rollout_equals_unsharded():
  import mujoco_warp_b200 as mjw

  mgr = mp.Manager()
  ret = mgr.dict()
  mp.spawn(_worker, args=(2, _free_port(), ret), nprocs=2, join=True)
  mjm = mjw.load_mjcf(H.HUMANOID)
  qpos, qvel, ctrl = H.make_states(mjm, NWORLD, seed=99)
  ref = _rollout(mjm, qpos, qvel, ctrl)
  assert ret["full"].shape == (NWORLD, mjm.nq + mjm.nv)
  np.testing.assert_array_equal(ret["full"], ref)
  assert ret["tmax"] == 2.0
