"""C-ABI boundary (CPU): libmyostep.so loads, exports every symbol include/myostep.h declares, the ctypes struct
mirrors the header, and compute entries fail loudly without a CUDA device (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, model_path


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "myostep.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(myo_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_exported():
    from myosuite_mjx_b200 import lib
    L = lib.lib()
    declared = _declared_symbols()
    assert len(declared) >= 18
    missing = [s for s in declared if not hasattr(L, s)]
    assert not missing, f"libmyostep.so lacks {missing}"
    assert sorted(lib.EXPORTS) == declared, "lib.EXPORTS must list exactly the header's entry points"


def test_no_oracle_symbols_in_product():
    """The product library must not link or embed the CPU oracle."""
    from myosuite_mjx_b200 import lib
    L = lib.lib()
    for s in ("orc_create", "orc_step", "orc_forward", "orc_rollout"):
        assert not hasattr(L, s)
    for root, _, files in os.walk(os.path.join(ROOT, "myosuite_mjx_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".h")):
                txt = open(os.path.join(root, f)).read()
                assert "import oracle" not in txt and "mjstep_oracle" not in txt and "oracle/" not in txt, f"{f} references the oracle"


def test_field_table():
    from myosuite_mjx_b200 import desc, lib
    L = lib.lib()
    names = [L.myo_field_name(i).decode() for i in range(len(desc.FIELD_NAMES))]
    assert [n.lower() for n in names] == [n.lower() for n in desc.FIELD_NAMES]
    assert L.myo_field_name(len(names)) == b""
    ints = {n for i, n in enumerate(names) if L.myo_field_is_int(i)}
    assert {"ncon", "nefc", "contact_geom", "contact_dim", "contact_efc_address", "efc_type", "solver_niter", "status"} <= ints
    assert all(L.myo_field_itemsize(i) == 4 for i in range(len(names)))


def test_desc_pack_matches_header(models):
    from myosuite_mjx_b200 import desc
    m = models("hand")
    d, keep = desc.pack(m)
    assert d.nq == m.nq and d.nv == m.nv and d.nu == m.nu and d.ntendon == m.ntendon
    assert abs(d.opt_timestep - 0.002) < 1e-12 and d.opt_gravity[2] == pytest.approx(-9.81)
    got = np.ctypeslib.as_array(d.body_mass, shape=(m.nbody,))
    assert np.array_equal(got, np.asarray(m.body_mass, dtype=np.float64))
    assert np.ctypeslib.as_array(d.actuator_trnid, shape=(m.nu, 2))[:, 1].tolist() == [-1] * m.nu


def test_fails_loudly_without_gpu(models):
    import torch
    if torch.cuda.is_available():
        pytest.skip("needs a box without CUDA")
    from myosuite_mjx_b200 import desc, lib
    L = lib.lib()
    d, keep = desc.pack(models("elbow"))
    h = ctypes.c_void_p()
    rc = L.myo_model_create(ctypes.byref(d), 0, ctypes.byref(h))
    assert rc == -3, "myo_model_create must return MYO_ERR_CUDA without a device"
    assert b"no CPU fallback" in L.myo_last_error()
    with pytest.raises(lib.MyoError):
        lib.check(rc)
    buf = (ctypes.c_float * 16)()
    assert L.myo_kat_quat(1, buf, buf, buf, buf, buf, None) == -3


def test_null_arguments_rejected():
    from myosuite_mjx_b200 import lib
    L = lib.lib()
    assert L.myo_model_create(None, 0, None) == -1
    assert L.myo_batch_bind(None, 0, None, 0) == -1
    assert L.myo_forward(None, None) == -1
    assert L.myo_step(None, 1, None) in (-1,)
    assert L.myo_field_count(None, 0) == 0


def test_fixtures_exist():
    for name in ("elbow", "finger", "hand", "hand_pose", "hand_sar", "legs"):
        assert os.path.exists(model_path(name))
