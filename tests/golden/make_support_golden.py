"""Golden vectors for get_state / set_state / contact_force / reset_data produced by EXECUTING the reference's own functions
(/root/reference/mujoco_warp/_src/support.py:354, :572, :711 and io.py:1458) under the numpy Warp interpreter.

For each case the reference steps the model twice from seeded states, then:
  * get_state(INTEGRATION) with and without a world mask,
  * contact_force for every live contact plus out-of-range ids, contact frame and world frame,
  * set_state of a seeded state vector under a mask,
  * reset_data under a mask, and reset_data of every world,
and all inputs / outputs go to `support_golden.npz`.  tests/test_oracle_support_golden.py checks the oracle (float64) and
tests/test_gpu_support.py the CUDA path against the file.

  python tests/golden/make_support_golden.py
"""

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, HERE)
sys.path.insert(0, ROOT)

import ref_runner as R  # noqa: E402
import reference_cases as RC  # noqa: E402

OUT = os.path.join(HERE, "support_golden.npz")
# case -> (reference_cases case giving model / states, cone)
CASES = {"humanoid_pyramidal": ("humanoid_step5", 0), "humanoid_elliptic": ("humanoid_elliptic_newton", 1), "equality": ("equality_friction_limits", 0),
         "box_spheres_pyramidal": ("box_and_spheres_on_plane", 0), "box_spheres_elliptic": ("box_and_spheres_on_plane", 1)}
STATE_FIELDS = "time qpos qvel act qacc_warmstart ctrl qfrc_applied xfrc_applied eq_active mocap_pos mocap_quat".split()
RESET_FIELDS = STATE_FIELDS + "qacc act_dot energy qM solver_niter ne nf nl nefc nacon".split()
CONTACT_FIELDS = "dist pos frame includemargin friction solref solreffriction solimp dim geom efc_address worldid type geomcollisionid".split()


def snap(d, fields, pre):
  out = {pre + f: np.array(getattr(d, f).numpy()) for f in fields}
  for f in CONTACT_FIELDS:
    out[pre + "contact_" + f] = np.array(getattr(d.contact, f).numpy())
  return out


def main():
  wp, ref = R.load_reference()
  types_mod = sys.modules["mujoco_warp._src.types"]
  sig = int(types_mod.State.INTEGRATION)
  store = {}
  for name, (base, cone) in CASES.items():
    c = RC.CASES[base]
    mjm, qpos, qvel, ctrl = RC.case_inputs(base)
    nworld = c["nworld"]
    m, d = R.make(mjm, nworld, c.get("nconmax", 24), c.get("njmax", 64))
    m.opt.cone = cone
    R.set_state(d, qpos, qvel, ctrl)
    rng = np.random.default_rng(100 + len(name))
    d.xfrc_applied.numpy()[...] = rng.normal(size=d.xfrc_applied.numpy().shape) * 0.1
    d.qfrc_applied.numpy()[...] = rng.normal(size=d.qfrc_applied.numpy().shape) * 0.1
    for _ in range(2):
      ref.step(m, d)
    rec = {"in_qpos": qpos, "in_qvel": qvel, "in_ctrl": ctrl, "in_xfrc_applied": np.array(d.xfrc_applied.numpy()),
           "in_qfrc_applied": np.array(d.qfrc_applied.numpy()), "sig": np.array(sig), "cone": np.array(cone)}  # fmt: skip
    rec.update(snap(d, RESET_FIELDS, "stepped_"))
    rec["stepped_efc_force"] = np.array(d.efc.force.numpy())
    # get_state
    nstate = 1 + mjm.nq + 3 * mjm.nv + mjm.na + mjm.nu + 6 * mjm.nbody + mjm.neq + 7 * mjm.nmocap
    active = np.arange(nworld) % 2 == 0
    st = wp.zeros((nworld, nstate), dtype=float)
    ref.get_state(m, d, st, sig)
    rec["state_all"] = np.array(st.numpy())
    st2 = wp.array(np.full((nworld, nstate), -7.0), dtype=float)
    ref.get_state(m, d, st2, sig, active=wp.array(active, dtype=bool))
    rec["state_masked"], rec["active"] = np.array(st2.numpy()), active
    # contact_force
    nacon = int(d.nacon.numpy()[0])
    ids = np.concatenate([np.arange(nacon), [nacon, nacon + 5]]).astype(np.int32)
    for to_world in (False, True):
      force = wp.array(np.full((len(ids), 6), -3.0), dtype=wp.spatial_vector)
      ref.contact_force(m, d, wp.array(ids, dtype=int), to_world, force)
      rec["force_world" if to_world else "force_contact"] = np.array(force.numpy()).reshape(len(ids), 6)
    rec["contact_ids"] = ids
    # set_state under a mask (no mocap bodies / at most one equality in these models: the reference's two cursor slips are inert)
    new_state = rng.normal(size=(nworld, nstate))
    if mjm.neq:
      col = 1 + mjm.nq + 3 * mjm.nv + mjm.na + mjm.nu + 6 * mjm.nbody
      new_state[:, col : col + mjm.neq] = rng.integers(0, 2, (nworld, mjm.neq))
    rec["new_state"] = new_state
    ref.set_state(m, d, wp.array(new_state, dtype=float), sig, active=wp.array(active, dtype=bool))
    rec.update(snap(d, STATE_FIELDS, "set_"))
    # reset_data under a mask, then everything
    reset = np.arange(nworld) % 3 != 1
    ref.reset_data(m, d, wp.array(reset, dtype=bool))
    rec["reset"] = reset
    rec.update(snap(d, RESET_FIELDS, "reset_"))
    ref.reset_data(m, d)
    rec.update(snap(d, RESET_FIELDS, "resetall_"))
    for k, v in rec.items():
      store[f"{name}/{k}"] = v
    print(f"{name:22s} nworld {nworld} nstate {nstate} nacon {nacon}  |force| max {np.abs(rec['force_world'][:nacon]).max() if nacon else 0.0:.3g}")
  np.savez_compressed(OUT, **store)
  print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
  main()
