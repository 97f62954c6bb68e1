"""Compile-time constants of a model (`mj_setConst` of MuJoCo 3.2.8; third-party, reached from the reference
through `mujoco.MjModel.from_xml_path`, myosuite/mjx/play.py:8).

A small float64 numpy/python evaluation of the position stage at `qpos0` (kinematics, subtree COM, spatial
tendons with wrapping, joint-space inertia) gives `tendon_length0`, `dof_invweight0`, `body_invweight0`,
`tendon_invweight0`, `actuator_acc0` and `stat_meaninertia`.  Host side, runs once per model; it is part of
the model compiler, not of the step path, and it is NOT the parity checker used by the tests.
"""
from __future__ import annotations

import math

import numpy as np

MINVAL = 1e-15
JNT_FREE, JNT_BALL, JNT_SLIDE, JNT_HINGE = 0, 1, 2, 3
WRAP_PULLEY, WRAP_SITE, WRAP_SPHERE, WRAP_CYLINDER = 2, 3, 4, 5


def _qmul(a, b):
    return np.array([
        a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
        a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
        a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
        a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


def _q2m(q):
    w, x, y, z = q
    return np.array([
        [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]])


# ---- 2-D circle wrapping (MuJoCo `mju_wrap` helpers; algorithm per the "Computation" chapter) ----------
def _intersect(p1, p2, p3, p4):
    det = (p4[1] - p3[1]) * (p2[0] - p1[0]) - (p4[0] - p3[0]) * (p2[1] - p1[1])
    if abs(det) < MINVAL:
        return False
    a = ((p4[0] - p3[0]) * (p1[1] - p3[1]) - (p4[1] - p3[1]) * (p1[0] - p3[0])) / det
    b = ((p2[0] - p1[0]) * (p1[1] - p3[1]) - (p2[1] - p1[1]) * (p1[0] - p3[0])) / det
    return 0 <= a <= 1 and 0 <= b <= 1


def _arc(p0, p1, ind):
    a = p0 / max(np.linalg.norm(p0), MINVAL)
    b = p1 / max(np.linalg.norm(p1), MINVAL)
    ang = math.acos(min(1.0, max(-1.0, float(a @ b))))
    cr = p0[1] * p1[0] - p0[0] * p1[1]
    # grazing path: cr is rounding noise and must not turn a ~0 arc into a ~2 pi one (same guard as the kernel)
    if abs(cr) > 1e-5 * np.linalg.norm(p0) * np.linalg.norm(p1) and ((cr > 0 and ind) or (cr < 0 and not ind)):
        ang = 2 * math.pi - ang
    return ang


def wrap_circle(d0, d1, sd, rad):
    """Returns (arc_length, t0, t1) or None."""
    sq0, sq1, sqr = d0 @ d0, d1 @ d1, rad * rad
    if sq0 < sqr or sq1 < sqr or rad < MINVAL:
        return None
    dif = d1 - d0
    dd = dif @ dif
    if dd < MINVAL:
        return None
    a = min(1.0, max(0.0, -(dif @ d0) / dd))
    near = d0 + a * dif
    if near @ near > sqr and (sd is None or sd @ near >= 0):
        return None
    s0, s1 = math.sqrt(sq0 - sqr), math.sqrt(sq1 - sqr)
    sols, good = [], []
    for i, sgn in enumerate((1.0, -1.0)):
        t0 = np.array([d0[0] * sqr + sgn * rad * d0[1] * s0, d0[1] * sqr - sgn * rad * d0[0] * s0]) / sq0
        t1 = np.array([d1[0] * sqr - sgn * rad * d1[1] * s1, d1[1] * sqr + sgn * rad * d1[0] * s1]) / sq1
        if sd is not None:
            t = t0 + t1
            t = t / max(np.linalg.norm(t), MINVAL)
            g = float(t @ sd)
        else:
            t = t0 - t1
            g = -float(t @ t)
        if _intersect(d0, t0, d1, t1):
            g = -10000.0
        sols.append((t0, t1))
        good.append(g)
    i = 0 if good[0] > good[1] else 1
    t0, t1 = sols[i]
    if _intersect(d0, t0, d1, t1):
        return None
    return rad * _arc(t0, t1, i), t0, t1


def wrap_inside(d0, d1, rad):
    l0, l1 = np.linalg.norm(d0), np.linalg.norm(d1)
    if l0 <= rad or l1 <= rad or rad < MINVAL or l0 < MINVAL or l1 < MINVAL:
        return None
    dif = d1 - d0
    dd = dif @ dif
    if dd > MINVAL:
        a = -(dif @ d0) / dd
        if 0 < a < 1:
            if np.linalg.norm(d0 + a * dif) <= rad:
                return None
    p = 0.5 * (d0 + d1)
    p = rad * p / max(np.linalg.norm(p), MINVAL)
    A, B = rad / l0, rad / l1
    cosG = (l0 * l0 + l1 * l1 - dd) / (2 * l0 * l1)
    if cosG < -1 + MINVAL:
        return None
    if cosG > 1 - MINVAL:
        return 0.0, p, p
    G = math.acos(cosG)
    z = 1 - 1e-7
    f = math.asin(A * z) + math.asin(B * z) - 2 * math.asin(z) + G
    if f > 0:
        return 0.0, p, p
    it = 0
    while it < 20 and abs(f) > 1e-6:
        df = (A / max(MINVAL, math.sqrt(1 - z * z * A * A)) + B / max(MINVAL, math.sqrt(1 - z * z * B * B))
              - 2 / max(MINVAL, math.sqrt(1 - z * z)))
        if df > -MINVAL:
            return 0.0, p, p
        z1 = z - f / df
        if z1 > z:
            return 0.0, p, p
        z = z1
        f = math.asin(A * z) + math.asin(B * z) - 2 * math.asin(z) + G
        if f > 1e-6:
            return 0.0, p, p
        it += 1
    if it >= 20:
        return 0.0, p, p
    if d0[0] * d1[1] - d0[1] * d1[0] > 0:
        vec, ang = d0 / l0, math.asin(z) - math.asin(A * z)
    else:
        vec, ang = d1 / l1, math.asin(z) - math.asin(B * z)
    c, s = math.cos(ang), math.sin(ang)
    p = rad * np.array([c * vec[0] - s * vec[1], s * vec[0] + c * vec[1]])
    return 0.0, p, p


def wrap_geom(x0, x1, gpos, gmat, radius, wtype, side):
    """3-D wrap of the segment x0->x1 over a sphere/cylinder.  Returns (wlen, w0, w1) or None."""
    p0 = gmat.T @ (x0 - gpos)
    p1 = gmat.T @ (x1 - gpos)
    if np.linalg.norm(p0) < MINVAL or np.linalg.norm(p1) < MINVAL:
        return None
    if wtype == WRAP_SPHERE:
        ax0 = p0 / np.linalg.norm(p0)
        nrm = np.cross(p0, p1)
        n = np.linalg.norm(nrm)
        if n < MINVAL:
            ax1 = np.ones(3)
            ax1[int(np.argmax(np.abs(ax0)))] = 0.0
            nrm = np.cross(ax0, ax1)
            nrm /= np.linalg.norm(nrm)
        else:
            nrm = nrm / n
        ax1 = np.cross(nrm, ax0)
        ax1 /= np.linalg.norm(ax1)
    else:
        ax0, ax1 = np.array([1.0, 0, 0]), np.array([0.0, 1, 0])
    d0 = np.array([p0 @ ax0, p0 @ ax1])
    d1 = np.array([p1 @ ax0, p1 @ ax1])
    sd, inside = None, False
    if side is not None:
        s = gmat.T @ (side - gpos)
        sd = np.array([s @ ax0, s @ ax1])
        # "inside" when the side site lies within the wrap radius (3-D norm in the geom frame, both wrap types)
        inside = np.linalg.norm(s) < radius
        sd = radius * sd / max(np.linalg.norm(sd), MINVAL)
    r = wrap_inside(d0, d1, radius) if inside else wrap_circle(d0, d1, sd, radius)
    if r is None:
        return None
    wlen, t0, t1 = r
    r0 = ax0 * t0[0] + ax1 * t0[1]
    r1 = ax0 * t1[0] + ax1 * t1[1]
    if wtype == WRAP_CYLINDER:
        L0 = math.hypot(p0[0] - r0[0], p0[1] - r0[1])
        L1 = math.hypot(p1[0] - r1[0], p1[1] - r1[1])
        tot = L0 + wlen + L1
        r0[2] = p0[2] + (p1[2] - p0[2]) * L0 / tot
        r1[2] = p0[2] + (p1[2] - p0[2]) * (L0 + wlen) / tot
        wlen = math.hypot(wlen, r1[2] - r0[2])
    return wlen, gmat @ r0 + gpos, gmat @ r1 + gpos


# ---- position stage at a given qpos (float64, numpy; compile-time only) -----------------------------------
def _rot_axis(axis, ang):
    s = math.sin(0.5 * ang)
    return np.array([math.cos(0.5 * ang), axis[0] * s, axis[1] * s, axis[2] * s])


def forward_pos(m, qpos):
    """Body/geom/site frames, spatial tendon lengths and dense moment arms, dense joint-space inertia."""
    nb, nv = m.nbody, m.nv
    xpos = np.zeros((nb, 3)); xquat = np.zeros((nb, 4)); xmat = np.zeros((nb, 3, 3))
    xquat[0, 0] = 1.0; xmat[0] = np.eye(3)
    xanchor = np.zeros((m.njnt, 3)); xaxis = np.zeros((m.njnt, 3))
    for b in range(1, nb):
        p = m.body_parentid[b]
        pos = xpos[p] + xmat[p] @ m.body_pos[b]
        quat = _qmul(xquat[p], m.body_quat[b])
        for j in range(m.body_jntadr[b], m.body_jntadr[b] + m.body_jntnum[b]) if m.body_jntnum[b] else ():
            qa = m.jnt_qposadr[j]
            t = m.jnt_type[j]
            if t == JNT_FREE:
                pos = np.array(qpos[qa:qa + 3], dtype=np.float64)
                quat = np.array(qpos[qa + 3:qa + 7], dtype=np.float64)
                quat /= np.linalg.norm(quat)
                xanchor[j] = pos
                xaxis[j] = np.array([0.0, 0, 1])
                continue
            R = _q2m(quat)
            xanchor[j] = R @ m.jnt_pos[j] + pos
            xaxis[j] = R @ m.jnt_axis[j]
            d = qpos[qa] - m.qpos0[qa]
            if t == JNT_SLIDE:
                pos = pos + xaxis[j] * d
            elif t == JNT_HINGE:
                quat = _qmul(quat, _rot_axis(m.jnt_axis[j], d))
                pos = xanchor[j] - _q2m(quat) @ m.jnt_pos[j]
            else:
                raise NotImplementedError("ball joint")
        quat = quat / np.linalg.norm(quat)
        xpos[b], xquat[b], xmat[b] = pos, quat, _q2m(quat)
    xipos = np.array([xpos[b] + xmat[b] @ m.body_ipos[b] for b in range(nb)])
    ximat = np.array([xmat[b] @ _q2m(m.body_iquat[b]) for b in range(nb)])
    gb = m.geom_bodyid
    geom_xpos = np.array([xpos[gb[g]] + xmat[gb[g]] @ m.geom_pos[g] for g in range(m.ngeom)]).reshape(-1, 3)
    geom_xmat = np.array([xmat[gb[g]] @ _q2m(m.geom_quat[g]) for g in range(m.ngeom)]).reshape(-1, 3, 3)
    sb = m.site_bodyid
    site_xpos = np.array([xpos[sb[s]] + xmat[sb[s]] @ m.site_pos[s] for s in range(m.nsite)]).reshape(-1, 3)

    def jac(point, body):
        jp = np.zeros((3, nv)); jr = np.zeros((3, nv))
        # last dof of the nearest ancestor with dofs
        b = body
        while b > 0 and m.body_dofnum[b] == 0:
            b = m.body_parentid[b]
        if b == 0:
            return jp, jr
        i = m.body_dofadr[b] + m.body_dofnum[b] - 1
        while i >= 0:
            j = m.dof_jntid[i]
            t = m.jnt_type[j]
            if t == JNT_HINGE:
                jr[:, i] = xaxis[j]
                jp[:, i] = np.cross(xaxis[j], point - xanchor[j])
            elif t == JNT_SLIDE:
                jp[:, i] = xaxis[j]
            elif t == JNT_FREE:
                k = i - m.jnt_dofadr[j]
                if k < 3:
                    jp[k, i] = 1.0
                else:
                    ax = xmat[m.jnt_bodyid[j]][:, k - 3]
                    jr[:, i] = ax
                    jp[:, i] = np.cross(ax, point - xpos[m.jnt_bodyid[j]])
            i = m.dof_parentid[i]
        return jp, jr

    # dense inertia: sum over bodies of J^T diag(m, I) J about each body's centre of mass
    M = np.zeros((nv, nv))
    for b in range(1, nb):
        jp, jr = jac(xipos[b], b)
        Iw = ximat[b] @ np.diag(m.body_inertia[b]) @ ximat[b].T
        M += m.body_mass[b] * jp.T @ jp + jr.T @ Iw @ jr
    M[np.diag_indices(nv)] += m.dof_armature

    # spatial tendons
    nt = m.ntendon
    ten_length = np.zeros(nt); ten_J = np.zeros((nt, nv))
    for t in range(nt):
        adr, num = m.tendon_adr[t], m.tendon_num[t]
        div = 1.0
        k = 0
        while k < num - 1:
            ty0 = m.wrap_type[adr + k]
            if ty0 == WRAP_PULLEY:
                div = m.wrap_prm[adr + k]
                k += 1
                continue
            ty1 = m.wrap_type[adr + k + 1]
            if ty1 == WRAP_PULLEY:
                k += 1
                continue
            s0 = m.wrap_objid[adr + k]
            if ty1 == WRAP_SITE:
                pts = [(site_xpos[s0], sb[s0]), (site_xpos[m.wrap_objid[adr + k + 1]], sb[m.wrap_objid[adr + k + 1]])]
                wlen = -1.0
                k += 1
            else:
                g = m.wrap_objid[adr + k + 1]
                s1 = m.wrap_objid[adr + k + 2]
                side = int(m.wrap_prm[adr + k + 1])
                r = wrap_geom(site_xpos[s0], site_xpos[s1], geom_xpos[g], geom_xmat[g], m.geom_size[g, 0], ty1,
                              site_xpos[side] if side >= 0 else None)
                if r is None:
                    pts = [(site_xpos[s0], sb[s0]), (site_xpos[s1], sb[s1])]
                    wlen = -1.0
                else:
                    wlen, w0, w1 = r
                    pts = [(site_xpos[s0], sb[s0]), (w0, gb[g]), (w1, gb[g]), (site_xpos[s1], sb[s1])]
                k += 2
            for a in range(0, len(pts) - 1, 1 if wlen < 0 else 2):
                (p0, b0), (p1, b1) = pts[a], pts[a + 1]
                dif = p1 - p0
                L = np.linalg.norm(dif)
                ten_length[t] += L / div
                if b0 != b1 and L > MINVAL:
                    j0, _ = jac(p0, b0)
                    j1, _ = jac(p1, b1)
                    ten_J[t] += (j1 - j0).T @ (dif / L) / div
            if wlen >= 0:
                ten_length[t] += wlen / div
    return dict(xpos=xpos, xquat=xquat, xmat=xmat, xipos=xipos, ximat=ximat, geom_xpos=geom_xpos, geom_xmat=geom_xmat,
                site_xpos=site_xpos, xanchor=xanchor, xaxis=xaxis, ten_length=ten_length, ten_J=ten_J, M=M, jac=jac)


def set_const(m):
    """Fill the qpos0-dependent model constants (see module docstring)."""
    nv = m.nv
    d = forward_pos(m, m.qpos0)
    M = d["M"]
    Minv = np.linalg.inv(M) if nv else np.zeros((0, 0))
    m.tendon_length0 = d["ten_length"].copy()
    ls = m.tendon_lengthspring.copy()
    for t in range(m.ntendon):
        if ls[t, 0] == -1 and ls[t, 1] == -1:
            ls[t] = m.tendon_length0[t]
    m.tendon_lengthspring = ls
    # dof_invweight0: diagonal of M^-1, averaged over the translational / rotational triple of a free joint
    inv = np.diag(Minv).copy() if nv else np.zeros(0)
    for j in range(m.njnt):
        if m.jnt_type[j] == JNT_FREE:
            a = m.jnt_dofadr[j]
            inv[a:a + 3] = inv[a:a + 3].mean()
            inv[a + 3:a + 6] = inv[a + 3:a + 6].mean()
    m.dof_invweight0 = inv
    m.dof_M0 = np.diag(M).copy() if nv else np.zeros(0)
    biw = np.zeros((m.nbody, 2))
    for b in range(1, m.nbody):
        if m.body_weldid[b] == 0:
            continue
        jp, jr = d["jac"](d["xipos"][b], b)
        biw[b, 0] = np.trace(jp @ Minv @ jp.T) / 3.0
        biw[b, 1] = np.trace(jr @ Minv @ jr.T) / 3.0
    m.body_invweight0 = biw
    J = d["ten_J"]
    m.tendon_invweight0 = np.einsum("ti,ij,tj->t", J, Minv, J) if m.ntendon else np.zeros(0)
    acc0 = np.zeros(m.nu)
    for i in range(m.nu):
        if m.actuator_trntype[i] == 3:
            mom = m.actuator_gear[i] * J[m.actuator_trnid[i]]
        else:
            mom = np.zeros(nv)
            mom[m.jnt_dofadr[m.actuator_trnid[i]]] = m.actuator_gear[i]
        acc0[i] = np.linalg.norm(Minv @ mom)
    m.actuator_acc0 = acc0
    m.stat_meaninertia = float(np.diag(M).mean()) if nv else 1.0
    m.M0_dense = M
