"""Shared helpers of the differential tests: run the oracle, load oracle state into the CUDA
backend (teacher forcing), compare.  The oracle is the reference's own code (oracle/_ref/oracle_step,
built by oracle/build_ref.py); it is the CHECKER here, never the thing under test."""
import json
import os
import subprocess
import tempfile

import numpy as np

from bradot_b200 import scenes, server

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE = os.path.join(ROOT, "oracle", "_ref", "oracle_step")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def have_oracle():
    return os.path.exists(ORACLE) and os.access(ORACLE, os.X_OK)


def run_oracle(sc, steps, dump_from=0, dump_every=1, threads=-1, keep=None):
    """Step `sc` with the reference; returns (timing dict, dump dict)."""
    d = tempfile.mkdtemp(prefix="bpc_oracle_")
    scn = os.path.join(d, "scene.bscn")
    dmp = keep or os.path.join(d, "out.bdmp")
    sc.write(scn)
    cmd = [ORACLE, scn, "--steps", str(steps), "--dump", dmp, "--dump-from", str(dump_from), "--dump-every", str(dump_every),
           "--threads", str(threads)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=1200)
    if r.returncode != 0:
        raise RuntimeError("oracle_step failed: %s" % r.stderr[-2000:])
    info = json.loads(r.stdout.strip().splitlines()[-1])
    dump = scenes.read_dump(dmp)
    os.remove(scn)
    if keep is None:
        os.remove(dmp)
        os.rmdir(d)
    return info, dump


def oracle_pairs(dump, step):
    """(ids, pf, cf) of an oracle step in the layout of bpc_pairs_write."""
    ids = dump[("pair_ids", step)].copy()
    pf = dump[("pair_f", step)][:, :3].copy()
    c = dump[("contacts", step)]
    n = len(ids)
    cf = np.zeros((n, 4, 16), np.float32)
    cf[:, :, 0:9] = c[:, :, 0:9]      # local_A, local_B, normal
    cf[:, :, 9] = c[:, :, 10]         # acc_normal
    cf[:, :, 10:13] = c[:, :, 11:14]  # acc_tangent
    cf[:, :, 13] = c[:, :, 14]        # acc_bias
    cf[:, :, 14] = c[:, :, 15]        # acc_bias_com
    cf[:, :, 15] = c[:, :, 20]        # used
    return ids, pf, cf


def load_state(ps, dump, step):
    """Teacher forcing: overwrite the GPU state with the oracle's state after `step`."""
    b = dump[("bodies", step)]
    ps.bodies_write(server.FIELD_TRANSFORM, b[:, 0:12])
    ps.bodies_write(server.FIELD_LINEAR_VELOCITY, b[:, 12:15])
    ps.bodies_write(server.FIELD_ANGULAR_VELOCITY, b[:, 15:18])
    ps.bodies_write(server.FIELD_ACTIVE, b[:, 18:19])
    ps.bodies_write(server.FIELD_STILL_TIME, b[:, 19:20])
    ps.bodies_write(server.FIELD_ITEM_BOX, dump[("items", step)])
    ids, pf, cf = oracle_pairs(dump, step)
    ps.pairs_write(ids, pf, cf)


def solve_order(dump, step):
    """The reference's sweep order of `step` as (kind, a, b) records for bpc_set_solve_order."""
    so = dump[("solve_order", step)]
    ids = dump[("pair_ids", step)]
    out = np.zeros((len(so), 3), np.int32)
    for i, (isl, kind, idx) in enumerate(so):
        assert kind == 0
        out[i] = (0, ids[idx, 0], ids[idx, 1])
    return out


def new_server(sc, dump=None):
    ps = server.PhysicsServerCUDA(0)
    meshes = server.meshes_from_dump(dump) if dump is not None else None
    ps.load_scene(sc, meshes)
    if dump is not None and ("mass_props", -1) in dump:
        ps.bodies_write(server.FIELD_MASS_PROPS, dump[("mass_props", -1)])
    return ps


def pair_keys(ids):
    return ids[:, 0].astype(np.int64) << 32 | ids[:, 1].astype(np.int64)


def rel_err(a, b, floor=1e-3):
    """max |a-b| / max(|b|, floor): relative where values are large, absolute-ish near zero."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor)))


def kinetic_energy_and_momentum(sc, bodies):
    """Translational kinetic energy and linear momentum of the dynamic bodies (mass from the scene)."""
    m = sc.bodies["mass"].astype(np.float64)
    dyn = sc.bodies["mode"] >= scenes.MODE_RIGID
    v = bodies[:, 12:15].astype(np.float64)
    y = bodies[:, 10].astype(np.float64)
    ke = 0.5 * np.sum(m[dyn] * np.sum(v[dyn] ** 2, axis=1))
    pe = np.sum(m[dyn] * 9.80665 * y[dyn])
    p = np.sum(m[dyn, None] * v[dyn], axis=0)
    return ke, pe, p
