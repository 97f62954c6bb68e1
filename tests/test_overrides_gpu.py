"""Per-world MODEL overrides (include/myostep.h MYO_F_GEOM_TYPE_OVR / _SIZE_OVR / _POS_OVR, MYO_F_ACT_GAINSCALE): what the
reference's envs rewrite in mjModel per episode -- the manipulated object's shape (reorient_sar_v0.py:207-245, incl. the box
of geom_type 6) and muscle weakness (base_v0.py:64-68) -- applied per WORLD on one compiled model.  Every world is compared
with a CPU oracle built from a model edited the same way."""
import copy

import numpy as np
import pytest

import parity

pytestmark = pytest.mark.gpu

# reorient_sar_v0.py:178-196: (geom_type, size) choices of Geometries8EnvV0
SHAPES = [(3, [0.013, 0.025, 0.025]), (3, [0.019, 0.040, 0.040]), (4, [0.011, 0.025, 0.025]), (4, [0.019, 0.040, 0.040]),
          (5, [0.013, 0.025, 0.025]), (5, [0.019, 0.040, 0.040]), (6, [0.017, 0.017, 0.017]), (6, [0.023, 0.023, 0.023])]


def _rbound(t, s):
    return {2: s[0], 3: s[0] + s[1], 4: max(s), 5: float(np.hypot(s[0], s[1])), 6: float(np.linalg.norm(s))}[t]


def _top_bot(t, s):   # reorient_sar_v0.py:214-239
    z = 1.3 * s[1] if t == 3 else (s[1] if t == 5 else s[2])
    return np.array([0, 0, z]), np.array([0, 0, -z])


def test_object_shape_per_world_matches_edited_models(models):
    import oracle
    import torch
    from myosuite_mjx_b200 import mjx
    m = models("hand_sar")
    gnames = [str(x) for x in m.names_geom]
    obj, top, bot = gnames.index("obj"), gnames.index("top"), gnames.index("bot")
    nw = 48
    qpos, qvel, act, ctrl = parity.random_state(m, nw, 11, qpos_spread=0.9)
    qpos[:, 23:] = 0.0                               # the object where the XML puts it, the fingers around (and partly inside) it
    model = mjx.put_model(m)
    d = mjx.put_data(model, dict(qpos=qpos, qvel=0 * qvel, act=act, ctrl=ctrl), nworld=nw,
                     fields=list(mjx.CORE_FIELDS) + ["ncon", "nefc", "contact_geom", "contact_dist", "contact_pos", "geom_xpos", "qacc", "efc_force",
                                                     "geom_type_ovr", "geom_size_ovr", "geom_pos_ovr"])
    shape = [SHAPES[w % 8] for w in range(nw)]
    for w, (t, s) in enumerate(shape):
        if w >= 40:
            continue          # the last worlds keep the compiled ellipsoid: overrides at their keep values
        d.geom_type_ovr[w, obj] = t
        d.geom_size_ovr[w, obj] = torch.tensor(s, device="cuda")
        tp, bp = _top_bot(t, s)
        d.geom_pos_ovr[w, top] = torch.tensor(tp, dtype=torch.float32, device="cuda")
        d.geom_pos_ovr[w, bot] = torch.tensor(bp, dtype=torch.float32, device="cuda")
    mjx.forward(model, d)
    torch.cuda.synchronize()
    assert int(d.status.abs().max()) == 0
    q, v, a = qpos, 0 * qvel, act
    with_obj, boxes = 0, 0
    for w in range(nw):
        mm = m
        if w < 40:
            t, s = shape[w]
            mm = copy.deepcopy(m)
            mm.geom_type[obj] = t; mm.geom_size[obj] = s; mm.geom_rbound[obj] = _rbound(t, s)
            mm.geom_pos[top], mm.geom_pos[bot] = _top_bot(t, s)
        o = oracle.Oracle(mm)
        o.reset(); o.set_state(q[w], v[w], a[w], ctrl[w], np.zeros(m.nv)); o.forward()
        n = o.ncon
        assert int(d.ncon[w, 0]) == n, f"world {w} ({shape[w]}): ncon {int(d.ncon[w, 0])} vs {n}"
        cg = np.asarray(o.contact_geom).reshape(-1, 2)
        assert np.array_equal(d.contact_geom[w, :n].cpu().numpy(), cg), f"world {w}: contact geoms"
        np.testing.assert_allclose(d.contact_dist[w, :n].cpu().numpy(), np.asarray(o.contact_dist), atol=5e-5)
        np.testing.assert_allclose(d.geom_xpos[w, [top, bot]].cpu().numpy(), np.asarray(o.geom_xpos).reshape(-1, 3)[[top, bot]], atol=2e-5)
        hit = (cg == obj).any()
        with_obj += int(hit); boxes += int(hit and w < 40 and shape[w][0] == 6)
    assert with_obj >= nw // 3 and boxes >= 4, f"too few worlds touch the object ({with_obj}, boxes {boxes}): the test poses need attention"


def test_box_rests_on_the_plane_with_four_contacts(tmp_path):
    """mjc_PlaneBox: four corner contacts carry the weight."""
    import torch
    import test_oracle_kat2 as K
    from myosuite_mjx_b200 import mjx
    m = K._compile(tmp_path, K.BALL.format(type="box", size="0.03 0.04 0.05"))
    model = mjx.put_model(m)
    q = np.tile(np.asarray(m.qpos0, dtype=np.float64), (8, 1)); q[:, 2] = 0.051
    d = mjx.put_data(model, dict(qpos=q), nworld=8, fields="all")
    for _ in range(6):
        mjx.step(model, d, 500)
    torch.cuda.synchronize()
    w = float(m.body_mass[1]) * K.G
    assert (d.ncon[:, 0] == 4).all() and (d.nefc[:, 0] == 16).all()
    np.testing.assert_allclose(d.efc_force[:, :16].double().sum(dim=1).cpu().numpy(), w, rtol=5e-4)
    assert float(d.qvel.abs().max()) < 1e-3


def test_gain_scale_per_world_is_sarcopenia(models):
    """MYO_F_ACT_GAINSCALE = 0.5 in some worlds equals a model whose actuator_gainprm[:, 2] is halved (base_v0.py:64-68)."""
    import torch
    from myosuite_mjx_b200 import mjx
    m = models("hand")
    nw = 64
    qpos, qvel, act, ctrl = parity.random_state(m, nw, 4, qpos_spread=0.4)
    weak = copy.deepcopy(m)
    weak.actuator_gainprm[:, 2] = 0.5 * weak.actuator_gainprm[:, 2]
    names = list(mjx.CORE_FIELDS) + ["qfrc_actuator"]
    a = mjx.put_data(mjx.put_model(m), dict(qpos=qpos, qvel=qvel, act=act, ctrl=ctrl), nworld=nw, fields=names + ["act_gainscale"])
    a.act_gainscale[::2] = 0.5
    b = mjx.put_data(mjx.put_model(weak), dict(qpos=qpos, qvel=qvel, act=act, ctrl=ctrl), nworld=nw, fields=names)
    c = mjx.put_data(mjx.put_model(m), dict(qpos=qpos, qvel=qvel, act=act, ctrl=ctrl), nworld=nw, fields=names)
    for x in (a, b, c):
        mjx.step(x.model, x, 5)
    torch.cuda.synchronize()
    for f in ("qpos", "qvel", "actuator_force", "qfrc_actuator"):
        assert torch.allclose(getattr(a, f)[::2], getattr(b, f)[::2], rtol=1e-6, atol=1e-7), f"{f}: weakened worlds"
        assert torch.equal(getattr(a, f)[1::2], getattr(c, f)[1::2]), f"{f}: untouched worlds"
