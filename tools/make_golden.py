"""Generates tests/golden/*.npz by running the REFERENCE (oracle/_ref, the reference's own
opende compiled in place, canonical-order mode) on small seeded scenes.  Run in the
container that has /root/reference mounted:  python tools/make_golden.py
Each file holds the body state after N steps plus the last step's pair list and contacts
of world 0; tests/scenes.py rebuilds the same scene from the same seed."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tests import scenes
from oracle.refharness import RefBatch

out = os.path.join(ROOT, "tests", "golden")
os.makedirs(out, exist_ok=True)
for name, (builder, steps) in scenes.GOLDEN.items():
    if os.environ.get('ONLY') and name != os.environ['ONLY']:
        continue
    sc, p, pd = builder()
    r = RefBatch(sc, p, mode=1)
    if pd is not None:
        r.set_hinge_pd(pd)
    r.step(steps)
    pos, quat, lvel, avel = r.body_state()
    rms, nc, ncj = r.stats()
    cg, cd = r.contacts(0)
    lam, lam_erp, cum = r.joint_state()
    np.savez_compressed(os.path.join(out, name + ".npz"), steps=steps, pos=pos, quat=quat, lvel=lvel,
                        avel=avel, pairs0=r.pairs(0), contact_geoms0=cg, contact_data0=cd,
                        n_contact_joints=ncj, rms=rms, flags=r.body_flags(), lam=lam, cum=cum)
    print(name, steps, "steps; contacts/world", ncj[:8], "file bytes",
          os.path.getsize(os.path.join(out, name + ".npz")))
