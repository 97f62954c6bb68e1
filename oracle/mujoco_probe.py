"""Probe for a real MuJoCo (TEST INFRASTRUCTURE, like the rest of oracle/).

BASELINE.md section 2 / SURVEY.md 8(c): the reference's physics is third-party `mujoco` 3.2.8, which is neither vendored in
/root/reference nor installed in this image, so parity of the physics step is UNPINNED.  If a `mujoco` python package ever
becomes importable (site-packages, or a driver-populated `baseline/_ref`) AND the MJCF files of the reference are reachable
(`MYO_MJCF_ROOT`, default /root/reference/myosuite), the harness switches to it by itself:
  * tests/test_mujoco_pin.py compares compiled constants and one mj_forward / mj_step field by field with the oracle;
  * bench.py's CPU arm times real `mj_step` (`cpu_baseline.kind = "mujoco"`).
Nothing here is imported by the product."""
from __future__ import annotations

import os
import sys
import threading

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MJCF = {  # model fixture -> MJCF path under the reference's myosuite/ tree
    "elbow": "simhive/myo_sim/elbow/myoelbow_1dof6muscles.xml",
    "finger": "simhive/myo_sim/finger/myofinger_v0.xml",
    "hand": "simhive/myo_sim/hand/myohand.xml",
    "hand_pose": "envs/myo/assets/hand/myohand_pose.xml",
    "hand_sar": "envs/myo/assets/hand/myohand_sar.xml",
    "legs": "simhive/myo_sim/leg/myolegs.xml",
}


def mujoco_module():
    """the `mujoco` module, or None"""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if os.path.isdir(ref) and ref not in sys.path:
        sys.path.append(ref)
    try:
        import mujoco  # noqa: F401
        return mujoco
    except Exception:
        return None


def mjcf_path(model_name):
    root = os.environ.get("MYO_MJCF_ROOT", "/root/reference/myosuite")
    p = os.path.join(root, MJCF.get(model_name, ""))
    return p if model_name in MJCF and os.path.exists(p) else None


def load(model_name):
    """(mujoco, MjModel) of a fixture's MJCF, or None when MuJoCo or the file is unavailable"""
    mj = mujoco_module()
    path = mjcf_path(model_name)
    if mj is None or path is None:
        return None
    return mj, mj.MjModel.from_xml_path(path)


class _Rollout:
    def __init__(self, mj, model, nworld, cores, qpos, seed):
        import numpy as np
        self.mj, self.model, self.cores = mj, model, cores
        self.data = [mj.MjData(model) for _ in range(nworld)]
        self.rng = [np.random.default_rng([seed, w]) for w in range(nworld)]
        self.t = 0
        for d, q in zip(self.data, qpos):
            d.qpos[:] = q

    def advance(self, nsteps, hold):
        def work(lo, hi):
            for w in range(lo, hi):
                d, t = self.data[w], self.t
                for k in range(nsteps):
                    if (t + k) % hold == 0:
                        d.ctrl[:] = self.rng[w].uniform(0, 1, self.model.nu)
                    self.mj.mj_step(self.model, d)   # releases the GIL
        n = len(self.data)
        chunk = (n + self.cores - 1) // self.cores
        th = [threading.Thread(target=work, args=(i, min(n, i + chunk))) for i in range(0, n, chunk)]
        [t.start() for t in th]; [t.join() for t in th]
        self.t += nsteps


def rollout_backend(model_name, nworld, cores, qpos, seed):
    got = load(model_name)
    if got is None:
        return None
    mj, model = got
    return _Rollout(mj, model, nworld, cores, qpos, seed)
