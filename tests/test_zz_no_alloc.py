"""myo_kinematics_noalloc (include/myostep.h): the allocation-free entry point
(no counterpart in the reference), against the existing entry point (myo_forward on a batch).

Collected LAST (file name).  The entry point's scratch is a fixed device array inside the
library that keeps the last call's contents, so the check is made on two calls in a row with different inputs and sizes
(the second one smaller, with a partial tile, over a region the first call left dirty).  The two paths run the same device
functions in the same order: the outputs are expected bit-identical; the assertion uses the 1e-6-of-scale bound of the
contact-record tests (tests/test_parity_fields_gpu.py) so that it reads as a tolerance."""
import ctypes

import numpy as np
import pytest

import parity

pytestmark = pytest.mark.gpu

OUTS = ["xpos", "xquat", "xmat", "site_xpos", "geom_xpos"]


def _existing(model, m, qpos, nw):
    import torch
    from myosuite_mjx_b200 import mjx
    d = mjx.put_data(model, dict(qpos=qpos), nworld=nw, fields="all")
    mjx.forward(model, d)
    torch.cuda.synchronize()
    return {k: d.storage(k).clone() for k in OUTS}   # the bound [k, nworld] buffers


def _noalloc(model, m, qpos, nw, skip=()):
    import torch
    from myosuite_mjx_b200 import lib
    L = lib.lib()
    width = {"xpos": 3 * m.nbody, "xquat": 4 * m.nbody, "xmat": 9 * m.nbody, "site_xpos": 3 * m.nsite, "geom_xpos": 3 * m.ngeom}
    q = torch.as_tensor(np.ascontiguousarray(qpos.T), dtype=torch.float32, device="cuda")   # [nq, nworld]
    out = {k: torch.full((width[k], nw), float("nan"), dtype=torch.float32, device="cuda") for k in OUTS if k not in skip}
    ptr = [ctypes.c_void_p(out[k].data_ptr()) if k in out else None for k in OUTS]
    torch.cuda.synchronize()
    rc = L.myo_kinematics_noalloc(model._h, nw, ctypes.c_void_p(q.data_ptr()), *ptr, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, L.myo_last_error().decode()
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("name", ["hand", "legs"])
def test_noalloc_kinematics_matches_forward_twice_in_a_row(models, name):
    from myosuite_mjx_b200 import mjx
    m = models(name)
    model = mjx.put_model(m)
    for call, (nw, seed, skip) in enumerate([(224, 11, ()), (77, 12, ("xmat",))]):   # the second call: fewer worlds, a partial tile, one output left out
        qpos = parity.random_state(m, nw, seed, qpos_spread=0.9)[0]
        ref = _existing(model, m, qpos, nw)
        got = _noalloc(model, m, qpos, nw, skip)
        for k, g in got.items():
            r = ref[k]
            assert g.shape == r.shape, (k, g.shape, r.shape)
            scale = max(float(r.abs().max()), 1e-12)
            err = float((g - r).abs().max())
            assert err <= 1e-6 * scale, f"{name} call {call}: {k} differs from myo_forward by {err:.3e} (scale {scale:.3e})"


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch(models):
    import torch
    from myosuite_mjx_b200 import lib, mjx
    m = models("hand")
    model = mjx.put_model(m)
    L = lib.lib()
    q = torch.zeros((m.nq, 32), dtype=torch.float32, device="cuda")
    x = torch.zeros((3 * m.nbody, 32), dtype=torch.float32, device="cuda")
    vp = ctypes.c_void_p
    rc = L.myo_kinematics_noalloc(model._h, 1 << 20, vp(q.data_ptr()), vp(x.data_ptr()), None, None, None, None, None)
    assert rc != 0 and "exceeds the fixed scratch" in L.myo_last_error().decode()
    assert L.myo_kinematics_noalloc(model._h, 32, vp(q.data_ptr()), None, None, None, None, None, None) != 0   # no output
    assert L.myo_kinematics_noalloc(None, 32, vp(q.data_ptr()), vp(x.data_ptr()), None, None, None, None, None) != 0
    assert L.myo_kinematics_noalloc(model._h, 32, vp(q.data_ptr()), vp(x.data_ptr()), None, None, None, None, None) == 0, L.myo_last_error().decode()
    torch.cuda.synchronize()
