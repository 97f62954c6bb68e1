"""Generates tests/golden/quat_kat.npz by IMPORTING the reference's numpy rotation helpers
(/root/reference/myosuite/utils/quat_math.py: mulQuat :15, axis_angle2quat :47, euler2quat :77, mat2quat :118,
quat2mat :160, rotVecQuat :209) on seeded inputs plus the hand-written cases of the reference's own test
(myosuite/tests/mjx/test_quat_math.py setUp).  Run in the build container only:
    python tests/golden/make_quat_kat.py [/root/reference]
The GPU box has no /root/reference: tests load the committed .npz."""
import importlib.util
import os
import sys

import numpy as np

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
spec = importlib.util.spec_from_file_location("ref_quat_math", os.path.join(REF, "myosuite", "utils", "quat_math.py"))
qm = importlib.util.module_from_spec(spec)
spec.loader.exec_module(qm)

rng = np.random.default_rng(20261104)
n = 256
qa = rng.normal(size=(n, 4)); qa /= np.linalg.norm(qa, axis=1, keepdims=True)
qb = rng.normal(size=(n, 4)); qb /= np.linalg.norm(qb, axis=1, keepdims=True)
# hand-written cases in the style of the reference test's setUp: identity, 90/180 degree turns about the axes
s = np.sqrt(0.5)
special = np.array([[1, 0, 0, 0], [s, s, 0, 0], [s, 0, s, 0], [s, 0, 0, s], [0, 1, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1],
                    [0.5, 0.5, 0.5, 0.5], [s, -s, 0, 0], [0.5, -0.5, 0.5, -0.5]], dtype=np.float64)
qa[:len(special)] = special
qb[:len(special)] = special[::-1]
vec = qb[:, 1:].copy()
axis = rng.normal(size=(n, 3)); axis /= np.linalg.norm(axis, axis=1, keepdims=True)
angle = rng.uniform(-np.pi, np.pi, n)
euler = rng.uniform(-1.5, 1.5, size=(n, 3))

out = dict(qa=qa, qb=qb, axis=axis, angle=angle, euler=euler,
           mul=np.array([qm.mulQuat(a, b) for a, b in zip(qa, qb)]),
           mat=np.array([qm.quat2mat(a) for a in qa]).reshape(n, 9),
           rot=np.array([qm.rotVecQuat(v, a) for v, a in zip(vec, qa)]),
           axisangle=np.array([qm.axis_angle2quat(a, t) for a, t in zip(axis, angle)]),
           mat2quat=np.array([qm.mat2quat(qm.quat2mat(a)) for a in qa]),
           euler2quat=np.array([qm.euler2quat(e) for e in euler]))
path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "quat_kat.npz")
np.savez_compressed(path, **out)
print("wrote", path, {k: v.shape for k, v in out.items()})
