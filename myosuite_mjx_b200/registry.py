"""Environment ids of the reference's gym registry for the BASELINE configs, with the registered kwargs restated
(reference: myosuite/envs/myo/myobase/__init__.py:123-137 elbow, :238-253 finger, :300-414 hand pose / ASL table,
:445-459 leg walk).  `make(env_id, nworld)` returns the batched, device-resident counterpart
(pipeline.PoseEnv / pipeline.WalkEnv): reset(rng) -> State, step(State, action) -> State."""
from __future__ import annotations

import numpy as np

# ASL digit poses 0..9 of the 23 hand joints (__init__.py:326-386), in the joint order of jnt_namesHand (:300-324)
HAND_JOINTS = ("pro_sup deviation flexion cmc_abduction cmc_flexion mp_flexion ip_flexion mcp2_flexion mcp2_abduction pm2_flexion "
               "md2_flexion mcp3_flexion mcp3_abduction pm3_flexion md3_flexion mcp4_flexion mcp4_abduction pm4_flexion md4_flexion "
               "mcp5_flexion mcp5_abduction pm5_flexion md5_flexion").split()
ASL_QPOS = np.array([r.split() for r in (
    "0 0 0 0.5624 0.28272 -0.75573 -1.309 1.30045 -0.006982 1.45492 0.998897 1.26466 0 1.40604 0.227795 1.07614 -0.020944 1.46103 0.06284 0.83263 -0.14399 1.571 1.38248",
    "0 0 0 0.0248 0.04536 -0.7854 -1.309 0.366605 0.010473 0.269258 0.111722 1.48459 0 1.45318 1.44532 1.44532 -0.204204 1.46103 1.44532 1.48459 -0.2618 1.47674 1.48459",
    "0 0 0 0.0248 0.04536 -0.7854 -1.13447 0.514973 0.010473 0.128305 0.111722 0.510575 0 0.37704 0.117825 1.44532 -0.204204 1.46103 1.44532 1.48459 -0.2618 1.47674 1.48459",
    "0 0 0 0.3384 0.25305 0.01569 -0.0262045 0.645885 0.010473 0.128305 0.111722 0.510575 0 0.37704 0.117825 1.571 -0.036652 1.52387 1.45318 1.40604 -0.068068 1.39033 1.571",
    "0 0 0 0.6392 -0.147495 -0.7854 -1.309 0.637158 0.010473 0.128305 0.111722 0.510575 0 0.37704 0.117825 0.306345 -0.010472 0.400605 0.133535 0.21994 -0.068068 0.274925 0.01571",
    "0 0 0 0.3384 0.25305 0.01569 -0.0262045 0.645885 0.010473 0.128305 0.111722 0.510575 0 0.37704 0.117825 0.306345 -0.010472 0.400605 0.133535 0.21994 -0.068068 0.274925 0.01571",
    "0 0 0 0.6392 -0.147495 -0.7854 -1.309 0.637158 0.010473 0.128305 0.111722 0.510575 0 0.37704 0.117825 0.306345 -0.010472 0.400605 0.133535 1.1861 -0.2618 1.35891 1.48459",
    "0 0 0 0.524 0.01569 -0.7854 -1.309 0.645885 -0.006982 0.128305 0.111722 0.510575 0 0.37704 0.117825 1.28036 -0.115192 1.52387 1.45318 0.432025 -0.068068 0.18852 0.149245",
    "0 0 0 0.428 0.22338 -0.7854 -1.309 0.645885 -0.006982 0.128305 0.194636 1.39033 0 1.08399 0.573415 0.667675 -0.020944 0 0.06284 0.432025 -0.068068 0.18852 0.149245",
    "0 0 0 0.5624 0.28272 -0.75573 -1.309 1.30045 -0.006982 1.45492 0.998897 0.39275 0 0.18852 0.227795 0.667675 -0.020944 0 0.06284 0.432025 -0.068068 0.18852 0.149245",
)], dtype=np.float64)

SPECS = {
    # id: (model fixture, env class, max_episode_steps, kwargs)
    "myoElbowPose1D6MRandom-v0": ("elbow", "pose", 100, dict(target_jnt_range={"r_elbow_flex": (0, 2.27)}, normalize_act=True, pose_thd=0.175,
                                                             reset_type="random")),
    "myoFingerPoseRandom-v0": ("finger", "pose", 100, dict(target_jnt_range={"IFadb": (-0.2, 0.2), "IFmcp": (-0.4, 1), "IFpip": (0.1, 1), "IFdip": (0.1, 1)},
                                                           normalize_act=True, reset_type="init")),
    "myoHandPoseRandom-v0": ("hand_pose", "pose", 100, dict(target_jnt_range={n: (ASL_QPOS[:, i].min(), ASL_QPOS[:, i].max()) for i, n in enumerate(HAND_JOINTS)},
                                                            normalize_act=True, pose_thd=0.7, reset_type="random", target_type="generate")),
    "myoLegWalk-v0": ("legs", "walk", 1000, dict(normalize_act=True, min_height=0.8, max_rot=0.8, hip_period=100, reset_type="init", target_x_vel=0.0,
                                                 target_y_vel=1.2, target_rot=None)),
    # Geometries8EnvV0: target orientation and object shape drawn per episode and per world
    "myoHandReorient8-v0": ("hand_sar", "reorient", 50, dict(normalize_act=True, n_frames=5)),
}
for _k in range(10):   # ASL evaluation envs, one per numeral (__init__.py:383-396)
    SPECS["myoHandPose%dFixed-v0" % _k] = ("hand_pose", "pose", 100, dict(target_jnt_value=ASL_QPOS[_k], normalize_act=True, pose_thd=0.7, reset_type="init",
                                                                          target_type="fixed"))


def ordered_targets(mj, spec_kwargs):
    """target_jnt_range / target_jnt_value in qpos order.  The reference indexes pose_err = target - qpos positionally
    (pose_v0.py:67-78,104), i.e. its dicts list every joint in qpos order; that is checked here."""
    kw = dict(spec_kwargs)
    rng = kw.pop("target_jnt_range", None)
    if rng is not None:
        names = list(mj.names_jnt)
        adr = [int(mj.jnt_qposadr[names.index(n)]) for n in rng]
        if adr != list(range(int(mj.nq))):
            raise ValueError("target_jnt_range must list every joint in qpos order")
        kw["target_jnt_range"] = np.array([rng[n] for n in rng], dtype=np.float32)
    return kw


def make(env_id, nworld=1, model=None, **overrides):
    """`model`: an mjcf.Model, a path to the env's MJCF, or None to load the committed fixture of that model."""
    from . import mjcf, pipeline
    fixture, kind, horizon, kwargs = SPECS[env_id]
    if model is None:
        import os
        model = mjcf.Model.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "models", fixture + ".npz"))
    elif isinstance(model, str):
        model = mjcf.from_xml_path(model)
    kw = ordered_targets(model, kwargs)
    if kind == "pose":
        kw.setdefault("target_type", "generate" if "target_jnt_range" in kw else "fixed")
    kw.update(horizon=horizon)
    kw.update(overrides)
    cls = {"pose": pipeline.PoseEnv, "walk": pipeline.WalkEnv, "reorient": pipeline.ReorientEnv}[kind]
    return cls(model, nworld=nworld, **kw)
