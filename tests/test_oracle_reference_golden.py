"""CPU oracle vs golden vectors produced by EXECUTING THE REFERENCE'S OWN CODE for the whole step path.

`tests/golden/reference_step.npz` was written by `tests/golden/make_reference_golden.py`, which imports the unmodified
source files under /root/reference/mujoco_warp and runs `put_model -> make_data -> forward / step` thread by thread in a
float64 numpy interpreter of the Warp API (`tests/golden/warp_interp.py`; mujoco enums from `mujoco_stub.py`).  This test
re-runs the oracle on the stored inputs: every Data field of the path (kinematics .. integrator, contacts, efc rows, solver
outputs, counts) must agree with the reference to 1e-9 (both sides float64).  That pins the oracle to the reference's
algorithm, row for row.  Contacts / efc rows are compared per world after a canonical sort (the reference appends them
through atomic counters: only the set is defined).
"""

import numpy as np
import pytest

from tests import helpers as H
from tests.golden import reference_cases as RC

TOL = 1e-9
TOL_LOOSE_SOLVER = 2e-2


@pytest.mark.parametrize("name", list(RC.CASES))
def test_oracle_matches_reference_execution(name):
  c = RC.CASES[name]
  g = RC.load_golden(name)
  assert g, f"case {name} missing from reference_step.npz: run tests/golden/make_reference_golden.py"
  mjm = RC.load_case_model(name)
  np.testing.assert_array_equal(RC.case_inputs(name)[1], g["in_qpos"])  # the fixture's inputs are the seeded ones
  om, od = H.oracle_data(mjm, g["in_qpos"], g["in_qvel"], g["in_ctrl"], c.get("nconmax", 24), c.get("njmax", 64))
  traj = []
  ctrl_traj = RC.case_ctrl_traj(name, mjm)
  for t in range(c.get("nstep", 1)):
    if ctrl_traj is not None:
      od.ctrl[:] = ctrl_traj[t]
    H.oracle_steps(om, od, c["mode"], 1, RC.case_opt(name))
    traj.append(od.qpos.copy())
  worst = RC.compare(name, g, od, verbose=False)
  loose = name in RC.LOOSE_SOLVER_CASES
  for f, e in worst.items():
    tol = TOL_LOOSE_SOLVER if loose and f in RC.SOLVER_FIELDS else RC.CASE_TOL.get(name, TOL)
    assert e < tol, (name, f, e)
  if not loose:
    assert H.max_err(np.stack(traj), g["traj_qpos"]) < RC.CASE_TOL.get(name, TOL)
  assert len(worst) > 40  # the comparison really covered the field list


def test_golden_covers_every_stage_output():
  g = RC.load_golden("humanoid_forward")
  for f in RC.REAL_FIELDS + RC.COUNTS + ["efc_" + x for x in RC.EFC_REAL + RC.EFC_INT] + ["contact_" + x for x in RC.CON_REAL + RC.CON_INT]:
    assert f in g, f
  assert int(g["nacon"][0]) > 0 and int(g["nefc"].max()) >= 32
