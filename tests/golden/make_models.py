"""Compiles the five BASELINE configs' MJCF files (read from the read-only reference checkout) into the
mjModel-shaped .npz files under myosuite_mjx_b200/models/ (shipped with the package: the GPU box has no /root/reference).  Run in the build container only:
    python tests/golden/make_models.py [/root/reference]
The GPU box has no /root/reference; tests, smoke() and bench.py load these fixtures instead."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from myosuite_mjx_b200 import mjcf  # noqa: E402

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
MODELS = {
    "elbow": "myosuite/envs/myo/assets/elbow/myoelbow_1dof6muscles.xml",    # config 1: myoElbowPose1D6MRandom-v0
    "finger": "myosuite/simhive/myo_sim/finger/myofinger_v0.xml",           # config 2: myoFingerPoseRandom-v0
    "hand": "myosuite/simhive/myo_sim/hand/myohand.xml",                    # config 3: myosuite/mjx/play.py:8
    "hand_pose": "myosuite/envs/myo/assets/hand/myohand_pose.xml",          # config 3: myoHandPoseRandom-v0
    "hand_sar": "myosuite/envs/myo/assets/hand/myohand_sar.xml",            # config 4: myoHandReorient8-v0
    "legs": "myosuite/simhive/myo_sim/leg/myolegs.xml",                     # config 5: myoLegWalk-v0
}
out = os.path.join(ROOT, "myosuite_mjx_b200", "models")
os.makedirs(out, exist_ok=True)
for name, rel in MODELS.items():
    m = mjcf.from_xml_path(os.path.join(REF, rel))
    m.save(os.path.join(out, name + ".npz"))
    print(name, "nq", m.nq, "nv", m.nv, "nu", m.nu, "ntendon", m.ntendon)
