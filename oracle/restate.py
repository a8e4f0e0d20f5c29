"""numpy restatement of the two simplest stages of the reference step, used ONLY by the tests to cross-check
the compiled oracle and the golden dumps (test infrastructure; nothing under bradot_b200/ imports this).

  * free-flight integration of a rigid body: BradotBody3D::integrate_forces / integrate_velocities /
    _update_transform_dependent (modules/bradot_physics_3d/bradot_body_3d.cpp:477-672, 674-738, 44-54)
  * the broadphase pair set as a function of the stored item boxes: closed-interval overlap
    (core/math/bvh_abb.h:217-225), never static-static (modules/bradot_physics_3d/bradot_broad_phase_3d_bvh.cpp:37-38)

All arithmetic is float32 in the reference's operation order.
"""
import numpy as np

F = np.float32


def _axis_angle(axis, angle):
    x, y, z = axis
    c, s = F(np.cos(angle, dtype=F)), F(np.sin(angle, dtype=F))
    t = F(1) - c
    m = np.zeros((3, 3), F)
    m[0, 0] = x * x + c * (F(1) - x * x)
    m[1, 1] = y * y + c * (F(1) - y * y)
    m[2, 2] = z * z + c * (F(1) - z * z)
    m[0, 1] = x * y * t - z * s
    m[1, 0] = x * y * t + z * s
    m[0, 2] = x * z * t + y * s
    m[2, 0] = x * z * t - y * s
    m[1, 2] = y * z * t - x * s
    m[2, 1] = y * z * t + x * s
    return m


def _orthonormalize(b):
    x, y, z = b[:, 0].copy(), b[:, 1].copy(), b[:, 2].copy()
    x = x / np.sqrt(np.dot(x, x), dtype=F)
    y = y - x * np.dot(x, y)
    y = y / np.sqrt(np.dot(y, y), dtype=F)
    z = z - x * np.dot(x, z) - y * np.dot(y, z)
    z = z / np.sqrt(np.dot(z, z), dtype=F)
    return np.stack([x, y, z], axis=1).astype(F)


def free_flight_step(xf, lin, ang, mass_props, dt, gravity=(0.0, -9.80665, 0.0), lin_damp=0.1, ang_damp=0.1):
    """One step of a contact-free rigid body.  xf: 12 floats (basis rows, origin); mass_props: the 16 floats of
    BPC_FIELD_MASS_PROPS / the oracle's mass_props dump.  Returns (xf, lin, ang)."""
    dt = F(dt)
    B = np.asarray(xf[:9], F).reshape(3, 3)
    o = np.asarray(xf[9:12], F)
    lin, ang = np.asarray(lin, F), np.asarray(ang, F)
    inv_mass = F(mass_props[0])
    iinv = np.asarray(mass_props[1:4], F)
    com_local = np.asarray(mass_props[4:7], F)
    axes = np.asarray(mass_props[7:16], F).reshape(3, 3)
    mass = F(1) / inv_mass
    # integrate_forces: damping first, then forces (no torque in free flight)
    force = np.asarray(gravity, F) * mass
    lin = lin * max(F(0), F(1) - dt * F(lin_damp))
    ang = ang * max(F(0), F(1) - dt * F(ang_damp))
    lin = lin + (force * inv_mass) * dt
    # integrate_velocities: rotate about the centre of mass, Gram-Schmidt, translate
    w = np.sqrt(np.dot(ang, ang), dtype=F)
    if not abs(w) < F(1e-5):
        R = _axis_angle(ang / w, w * dt)
        o = o + ((np.eye(3, dtype=F) - R) @ B) @ com_local
        B = _orthonormalize((R @ B).astype(F))
    o = o + lin * dt
    return np.concatenate([B.reshape(-1), o]).astype(F), lin.astype(F), ang.astype(F)


def overlap_pairs(stored_min, stored_max, static, layer=None, mask=None):
    """All (a, b), a < b, whose stored boxes overlap (closed intervals), excluding static-static."""
    n = len(stored_min)
    out = set()
    for a in range(n):
        ov = np.all(stored_min[a] <= stored_max[a + 1:], axis=1) & np.all(stored_min[a + 1:] <= stored_max[a], axis=1)
        for b in np.nonzero(ov)[0] + a + 1:
            if static[a] and static[b]:
                continue
            if layer is not None and not ((layer[a] & mask[b]) or (layer[b] & mask[a])):
                continue
            out.add((a, int(b)))
    return out
