"""Developer tool (GPU box): throughput of the fused env layer through its public API (registry.make(...).step), 8192 worlds,
random actions resident on the device.  Prints one JSON line per env id: physics env-steps/s (= worlds x frame_skip x env
steps / time, the BASELINE metric's unit), env-level steps/s, kernel launches per env step, episodes restarted.
usage: python tools/env_bench.py [nworld] [timed env steps]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from myosuite_mjx_b200 import lib, registry   # noqa: E402

nw = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
nsteps = int(sys.argv[2]) if len(sys.argv) > 2 else 100
for env_id in ("myoElbowPose1D6MRandom-v0", "myoFingerPoseRandom-v0", "myoHandPoseRandom-v0", "myoHandReorient8-v0", "myoLegWalk-v0"):
    env = registry.make(env_id, nworld=nw)
    st = env.reset(1234)
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    acts = [torch.rand((nw, env.action_size), generator=g, device="cuda") * 2 - 1 for _ in range(8)]
    for i in range(20):
        st = env.step(st, acts[i % 8])
    torch.cuda.synchronize()
    l0 = lib.lib().myo_launch_count()
    ep0 = int(env.env.episode.sum())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(nsteps):
        st = env.step(st, acts[i % 8])
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    fs = env.cfg["frame_skip"]
    print(json.dumps({"env": env_id, "nworld": nw, "frame_skip": fs, "env_steps": nsteps, "ms_per_env_step": ms / nsteps,
                      "physics_env_steps_per_s": nw * fs * nsteps / (ms * 1e-3), "env_level_steps_per_s": nw * nsteps / (ms * 1e-3),
                      "launches_per_env_step": (lib.lib().myo_launch_count() - l0) / nsteps, "episodes_restarted": int(env.env.episode.sum()) - ep0,
                      "obs_dim": int(st.obs.shape[1]), "finite": bool(torch.isfinite(st.obs).all())}), flush=True)
    del env, st
