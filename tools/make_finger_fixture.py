"""GPU box: fills the finger fixture's muscle length ranges with the compile-time simulation (lengthrange.py) and
writes the completed model to gpurun_out/finger.npz (copied to myosuite_mjx_b200/models/finger.npz afterwards)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from myosuite_mjx_b200 import mjcf, lengthrange
m = mjcf.Model.load(os.path.join(ROOT, "myosuite_mjx_b200/models/finger.npz"))
lr = np.asarray(m.actuator_lengthrange, dtype=np.float64).reshape(-1, 2).copy(); lr[:] = 0   # recompute from scratch
m.actuator_lengthrange = lr
ids, rng, spread = lengthrange.set_length_range(m)
print("actuators", ids, "\nlengthrange\n", rng, "\nresidual spread\n", spread)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
m.save(os.path.join(ROOT, "gpurun_out", "finger.npz"))
