"""Golden vectors for the MJCF compiler's derived constants, produced by EXECUTING the reference's own `set_const_fixed` /
`set_const_0` (reference io.py:2197, :2222; kernels :1750-2190) under the numpy Warp interpreter.

`mujoco_warp_b200/mjcf.py` has to compute `body_subtreemass`, `dof_invweight0`, `body_invweight0`, `stat.meaninertia` and
`actuator_acc0` itself (MuJoCo's compiler is not installed); these feed every efc_D / aref / solver tolerance.  The reference
recomputes exactly these fields on the device from masses, inertias and qpos0, so its outputs pin the compiler:

  python tests/golden/make_setconst_golden.py          # writes tests/golden/setconst_golden.npz (needs /root/reference)

`tests/test_mjcf_constants_golden.py` compares the committed file with what `mjcf.py` computes (CPU test, no reference needed).
"""

from __future__ import annotations

import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_runner as R  # noqa: E402
import reference_cases as RC  # noqa: E402
from tests import models as TM  # noqa: E402

OUT = os.path.join(HERE, "setconst_golden.npz")
MODELS = ["humanoid", "panda", "apollo", "aloha"] + list(TM.XML)
FIELDS = ("body_subtreemass", "dof_invweight0", "body_invweight0", "actuator_acc0")


def run(key):
  wp, ref = R.load_reference()
  io_mod = sys.modules["mujoco_warp._src.io"]
  mjm = RC.load_model(key)
  from mujoco_warp_b200 import constants as C

  mjm.opt.disableflags |= int(C.DisableBit.NATIVECCD)  # the reference's put_model refuses box-box margins otherwise (io.py:386-395); no effect on the constants
  m, d = R.make(mjm, 1, 8, 16)
  for f in FIELDS:  # poison what the reference is about to recompute: the outputs cannot be the compiler's own numbers handed back
    getattr(m, f).numpy()[...] = np.nan
  m.stat.meaninertia.numpy()[...] = np.nan
  io_mod.set_const_fixed(m, d)
  io_mod.set_const_0(m, d)
  out = {f: np.array(getattr(m, f).numpy())[0] for f in FIELDS}
  out["meaninertia"] = np.array(m.stat.meaninertia.numpy()).reshape(-1)[:1]
  return mjm, out


def main():
  store = {}
  for key in MODELS:
    t = time.time()
    mjm, out = run(key)
    worst = 0.0
    for f, ref in out.items():
      mine = np.asarray(mjm.stat.meaninertia if f == "meaninertia" else getattr(mjm, f), dtype=np.float64).reshape(ref.shape)
      err = float(np.max(np.abs(mine - ref) / (1.0 + np.abs(ref)))) if ref.size else 0.0
      worst = max(worst, err)
      store[f"{key}/{f}"] = ref
    print(f"{key:32s} {time.time() - t:6.1f}s  nbody {mjm.nbody:3d} nv {mjm.nv:3d}  mjcf.py vs reference set_const: worst {worst:.2e}")
  np.savez_compressed(OUT, **store)
  print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
  main()
