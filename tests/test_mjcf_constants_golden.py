"""The MJCF compiler's derived constants (`mjcf._set_const`: body_subtreemass, dof_invweight0, body_invweight0, stat.meaninertia,
actuator_acc0) against the reference's OWN recomputation of them: `tests/golden/setconst_golden.npz` holds the outputs of the
reference's `set_const_fixed` + `set_const_0` (io.py:2197, :2222) executed from its unmodified source on every benchmark and test
model (`tests/golden/make_setconst_golden.py`).  These constants scale every efc_D / aref and the solver tolerance, and MuJoCo's own
compiler is not available to compare with.  Bar: 1e-10 relative (measured 5e-15)."""

import os

import numpy as np
import pytest

from tests.golden import reference_cases as RC

NPZ = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "setconst_golden.npz")
MODELS = ["humanoid", "panda", "apollo", "aloha", "equality_friction_limits", "box_and_spheres_on_plane", "connect_weld", "springs", "gravcomp",
          "site_equalities", "no_constraints_no_actuators", "rk4_actuators", "box_pairs", "convex_pairs", "mesh_pairs"]  # fmt: skip
FIELDS = ("body_subtreemass", "dof_invweight0", "body_invweight0", "actuator_acc0", "meaninertia")


@pytest.mark.parametrize("key", MODELS)
def test_compiler_constants_match_reference_set_const(key):
  mjm = RC.load_model(key)
  with np.load(NPZ) as z:
    for f in FIELDS:
      ref = z[f"{key}/{f}"]
      assert np.isfinite(ref).all(), f  # the generator poisons the fields with NaN before the reference recomputes them
      mine = np.asarray(mjm.stat.meaninertia if f == "meaninertia" else getattr(mjm, f), dtype=np.float64).reshape(ref.shape)
      if f != "actuator_acc0":
        assert ref.size > 0, f
      err = float(np.max(np.abs(mine - ref) / (1.0 + np.abs(ref)))) if ref.size else 0.0
      assert err < 1e-10, (key, f, err)
  # the interesting entries are not trivially zero
  assert np.asarray(mjm.dof_invweight0).max() > 0 and np.asarray(mjm.body_invweight0)[1:].max() > 0
