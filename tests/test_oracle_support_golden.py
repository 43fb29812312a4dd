"""Oracle restatements of get_state / set_state / contact_force / reset_data (oracle/oracle.py) vs outputs of the REFERENCE's
own functions (support.py:354/:572/:711, io.py:1458) executed by tests/golden/make_support_golden.py.

The reference's Data after two steps (state, contacts, efc_force, counts) is loaded into the oracle's Data, then the same call
sequence runs on both sides; every output must agree exactly (these are copies, sums and a 3x3 rotation in float64)."""

import numpy as np
import pytest

import mujoco_warp_b200 as mjw
from oracle import oracle as O
from tests import helpers as H
from tests.golden import reference_cases as RC

TOL = 1e-12


def _oracle_case(name):
  rec = H.load_support_case(name)
  mjm = RC.load_model(RC.CASES[H.SUPPORT_CASES[name]]["model"])
  c = RC.CASES[H.SUPPORT_CASES[name]]
  om = O.OracleModel(mjm)
  od = O.make_data(om, nworld=c["nworld"], nconmax=c.get("nconmax", 24), njmax=c.get("njmax", 64))

  def assign(f, a):
    dst = getattr(od, f)
    dst.reshape(-1)[:] = np.asarray(a).astype(dst.dtype).reshape(-1)

  H.support_fill(rec, "stepped_", H.SUPPORT_RESET_FIELDS + H.SUPPORT_CONTACT_FIELDS + ["efc_force"], assign)
  return rec, om, od


def _check(od, rec, prefix, fields):
  for f in fields:
    if prefix + f not in rec:
      continue
    want, got = np.asarray(rec[prefix + f], dtype=np.float64).reshape(-1), np.asarray(getattr(od, f), dtype=np.float64).reshape(-1)
    assert want.size == got.size, (f, want.size, got.size)
    assert np.abs(want - got).max(initial=0.0) <= TOL, (prefix, f, np.abs(want - got).max())


@pytest.mark.parametrize("name", list(H.SUPPORT_CASES))
def test_support_functions_match_reference_execution(name):
  rec, om, od = _oracle_case(name)
  sig, cone, active = int(rec["sig"]), int(rec["cone"]), rec["active"]
  assert sig == int(mjw.State.INTEGRATION)
  np.testing.assert_allclose(O.get_state(om, od, sig), rec["state_all"], atol=TOL, rtol=0)
  masked = np.full_like(rec["state_masked"], -7.0)
  O.get_state(om, od, sig, active=active, state=masked)
  np.testing.assert_allclose(masked, rec["state_masked"], atol=TOL, rtol=0)
  for key, to_world in (("force_contact", False), ("force_world", True)):
    force = np.full_like(rec[key], -3.0)
    O.contact_force(om, od, rec["contact_ids"], to_world, cone=cone, force=force)
    np.testing.assert_allclose(force, rec[key], atol=1e-10, rtol=0, err_msg=key)
  O.set_state(om, od, rec["new_state"], sig, active=active)
  _check(od, rec, "set_", H.SUPPORT_STATE_FIELDS)
  O.reset_data(om, od, rec["reset"])
  _check(od, rec, "reset_", H.SUPPORT_RESET_FIELDS + H.SUPPORT_CONTACT_FIELDS)
  O.reset_data(om, od)
  _check(od, rec, "resetall_", H.SUPPORT_RESET_FIELDS + H.SUPPORT_CONTACT_FIELDS)


def test_state_size_and_signature_validation():
  mjm = mjw.load_mjcf(H.HUMANOID)
  m = mjw.put_model(mjm, device="cpu")
  d = mjw.make_data(mjm, nworld=2, device="cpu")
  assert mjw.state_size(m, d, mjw.State.INTEGRATION) == 1 + mjm.nq + 3 * mjm.nv + mjm.na + mjm.nu + 6 * mjm.nbody
  assert mjw.state_size(m, d, mjw.State.PHYSICS) == mjm.nq + mjm.nv + mjm.na
  with pytest.raises(ValueError, match="invalid state signature"):
    mjw.state_size(m, d, 1 << 13)
