"""Developer script (GPU box): tendon path points, CUDA path vs oracle, per model."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import parity
from myosuite_mjx_b200 import mjcf
for name in sys.argv[1:] or ["finger", "hand", "legs", "hand_sar", "elbow"]:
    m = mjcf.Model.load(os.path.join(ROOT, "myosuite_mjx_b200/models", name + ".npz"))
    nw = 32
    qpos, qvel, act, ctrl = parity.random_state(m, nw, 1234, qpos_spread=0.9)
    ref, ex = parity.oracle_forward(m, qpos, qvel, act, ctrl, ["ten_length"])
    model, d, got = parity.gpu_forward(m, qpos, qvel, act, ctrl, nconmax=32, njmax=160)
    bad = [w for w in range(nw) if not np.array_equal(got.ten_wrapnum[w], ex["ten_wrapnum"][w])]
    print(f"== {name}: worlds with different ten_wrapnum: {bad}")
    for w in bad[:3]:
        print("  world", w, "gpu num", got.ten_wrapnum[w], "adr", got.ten_wrapadr[w])
        print("          orc num", ex["ten_wrapnum"][w], "adr", ex["ten_wrapadr"][w])
        n = int(ex["ten_wrapadr"][w][-1] + ex["ten_wrapnum"][w][-1])
        print("   gpu obj", got.wrap_obj.reshape(nw, -1)[w, :n + 4])
        print("   orc obj", ex["wrap_obj"][w][:n + 4])
        print("   ten_length err", np.abs(got.ten_length[w] - ref["ten_length"][w]))
