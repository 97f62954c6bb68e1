"""MJCF -> flat `mjModel`-shaped arrays, for the MJCF subset the MyoSuite models use.

Stands in for `mujoco.MjModel.from_xml_path` (reference call sites:
myosuite/mjx/play.py:8, myosuite/mjx/myodm_v0.py:43, myosuite/physics/mj_sim_scene.py:40-44),
which is third-party C that is not installed here.  Field names, enum values and index orders follow
MuJoCo 3.2.8's `mjModel` so that a real `mujoco.MjModel` can be passed to `put_model` instead.

Host-side, float64, numpy only.  Nothing here runs per step.
"""
from __future__ import annotations

import math
import os
import struct
import xml.etree.ElementTree as ET

import numpy as np

from . import setconst

# ---- enums (values as in mujoco/mjmodel.h) -------------------------------------------------------
JNT_FREE, JNT_BALL, JNT_SLIDE, JNT_HINGE = 0, 1, 2, 3
GEOM_PLANE, GEOM_HFIELD, GEOM_SPHERE, GEOM_CAPSULE, GEOM_ELLIPSOID, GEOM_CYLINDER, GEOM_BOX, GEOM_MESH = range(8)
GEOM_NAMES = {"plane": 0, "hfield": 1, "sphere": 2, "capsule": 3, "ellipsoid": 4, "cylinder": 5, "box": 6, "mesh": 7}
WRAP_NONE, WRAP_JOINT, WRAP_PULLEY, WRAP_SITE, WRAP_SPHERE, WRAP_CYLINDER = range(6)
DYN_NONE, DYN_INTEGRATOR, DYN_FILTER, DYN_FILTEREXACT, DYN_MUSCLE = 0, 1, 2, 3, 4
GAIN_FIXED, GAIN_AFFINE, GAIN_MUSCLE = 0, 1, 2
BIAS_NONE, BIAS_AFFINE, BIAS_MUSCLE = 0, 1, 2
TRN_JOINT, TRN_TENDON = 0, 3
EQ_JOINT = 2
NPRM = 10
MINVAL = 1e-15


class MjcfError(NotImplementedError):
    """Unsupported MJCF feature (raised at compile time, like MJX's put_model)."""


# ---- small math ------------------------------------------------------------------------------------
def _floats(s, n=None):
    v = np.array([float(x) for x in s.replace(",", " ").split()], dtype=np.float64)
    if n is not None and v.size != n:
        raise MjcfError(f"expected {n} numbers, got '{s}'")
    return v


def quat_mul(a, b):
    return np.array([
        a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
        a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
        a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
        a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


def quat_from_axis_angle(axis, ang):
    s = math.sin(0.5 * ang)
    return np.array([math.cos(0.5 * ang), axis[0] * s, axis[1] * s, axis[2] * s])


def quat_to_mat(q):
    w, x, y, z = q
    return np.array([
        [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z]])


def mat_to_quat(R):
    # Shepperd's method on a proper rotation matrix
    t = np.trace(R)
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        q = np.array([0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s])
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = math.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = np.array([(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s])
    elif R[1, 1] > R[2, 2]:
        s = math.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = np.array([(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s])
    else:
        s = math.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = np.array([(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s])
    q /= np.linalg.norm(q)
    if q[0] < 0:
        q = -q
    return q


def quat_z_to_vec(vec):
    """Quaternion rotating +z onto `vec` (used by fromto / zaxis)."""
    v = np.asarray(vec, dtype=np.float64)
    n = np.linalg.norm(v)
    if n < 1e-10:
        return np.array([1.0, 0, 0, 0])
    v = v / n
    axis = np.cross([0.0, 0.0, 1.0], v)
    s = np.linalg.norm(axis)
    if s < 1e-10:
        axis = np.array([1.0, 0.0, 0.0])
    else:
        axis = axis / s
    ang = math.atan2(s, v[2])
    return quat_from_axis_angle(axis, ang)


def principal_axes(I):
    """Symmetric 3x3 -> (eigenvalues descending, quaternion of the right-handed eigenframe)."""
    w, V = np.linalg.eigh(I)
    order = np.argsort(-w)
    w = w[order]
    V = V[:, order]
    if np.linalg.det(V) < 0:
        V[:, 2] = -V[:, 2]
    return w, mat_to_quat(V)


# ---- mesh mass properties ----------------------------------------------------------------------------
def load_stl(path, scale):
    with open(path, "rb") as f:
        data = f.read()
    ntri = struct.unpack_from("<I", data, 80)[0]
    if 84 + 50 * ntri != len(data):
        raise MjcfError(f"{path}: not a binary STL")
    rec = np.frombuffer(data, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=ntri, offset=84)
    return rec["v"].astype(np.float64) * np.asarray(scale, dtype=np.float64)


def mesh_mass_properties(tri):
    """Unit-density volume, centre of mass and inertia about the centre of mass of a closed triangle mesh
    (signed tetrahedra against the origin)."""
    a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]
    vol6 = np.einsum("ij,ij->i", a, np.cross(b, c))
    vol = vol6.sum() / 6.0
    sign = 1.0 if vol >= 0 else -1.0
    com = ((a + b + c) * vol6[:, None]).sum(0) / (24.0 * vol)
    # second moments  integral x x^T dV over each tetrahedron (origin, a, b, c)
    S = np.zeros((3, 3))
    for p, q in ((a, a), (b, b), (c, c)):
        S += np.einsum("i,ij,ik->jk", vol6, p, q) * 2
    for p, q in ((a, b), (a, c), (b, c)):
        S += np.einsum("i,ij,ik->jk", vol6, p, q) + np.einsum("i,ij,ik->jk", vol6, q, p)
    S /= 120.0
    vol *= sign
    S *= sign
    S -= vol * np.outer(com, com)
    I = np.trace(S) * np.eye(3) - S
    return vol, com, I


def hull_vertices(points):
    """Vertices of the convex hull of a point set (unique, in the hull's vertex order)."""
    from scipy.spatial import ConvexHull
    pts = np.unique(np.asarray(points, dtype=np.float64), axis=0)
    return pts[ConvexHull(pts).vertices]


def hull_triangles(points):
    """Outward-oriented triangles of the convex hull of a point set."""
    from scipy.spatial import ConvexHull
    pts = np.unique(np.asarray(points, dtype=np.float64), axis=0)
    h = ConvexHull(pts)
    tri = pts[h.simplices]
    n = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    flip = np.einsum("ij,ij->i", n, h.equations[:, :3]) < 0
    tri[flip] = tri[flip][:, ::-1]
    return tri


# ---- XML loading ---------------------------------------------------------------------------------
def _load_tree(path, main_dir):
    root = ET.parse(path).getroot()
    _expand_includes(root, os.path.dirname(os.path.abspath(path)), main_dir)
    return root


def _expand_includes(elem, cur_dir, main_dir):
    i = 0
    while i < len(elem):
        child = elem[i]
        if child.tag == "include":
            f = child.attrib["file"]
            cand = [os.path.join(cur_dir, f), os.path.join(main_dir, f)]
            p = next((c for c in cand if os.path.exists(c)), None)
            if p is None:
                raise MjcfError(f"include not found: {f}")
            sub = ET.parse(p).getroot()
            _expand_includes(sub, os.path.dirname(os.path.abspath(p)), main_dir)
            elem.remove(child)
            for k, c in enumerate(list(sub)):
                elem.insert(i + k, c)
            i += len(sub)
        else:
            _expand_includes(child, cur_dir, main_dir)
            i += 1


# ---- defaults ----------------------------------------------------------------------------------------
_ACT_TAGS = ("general", "muscle", "motor", "position", "velocity")


def _actuator_global_default():
    d = dict(dyntype=DYN_NONE, gaintype=GAIN_FIXED, biastype=BIAS_NONE,
             dynprm=np.zeros(NPRM), gainprm=np.zeros(NPRM), biasprm=np.zeros(NPRM),
             ctrllimited=None, ctrlrange=np.zeros(2), forcelimited=None, forcerange=np.zeros(2),
             actlimited=None, actrange=np.zeros(2), lengthrange=np.zeros(2), gear=np.array([1.0, 0, 0, 0, 0, 0]))
    d["dynprm"][0] = 1.0
    d["gainprm"][0] = 1.0
    return d


def _copy_act(d):
    return {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in d.items()}


def _tristate(s):
    return {"true": True, "false": False, "auto": None}[s]


def _apply_actuator(tag, a, base):
    """Fold one <general>/<muscle>/<position>/<motor> element (or default) into an actuator record."""
    d = _copy_act(base)
    for key in ("ctrlrange", "forcerange", "actrange", "lengthrange"):
        if key in a:
            d[key] = _floats(a[key], 2)
    for key in ("ctrllimited", "forcelimited", "actlimited"):
        if key in a:
            d[key] = _tristate(a[key])
    if "gear" in a:
        g = _floats(a["gear"])
        d["gear"] = np.concatenate([g, np.zeros(6 - g.size)])
    if tag == "general":
        names = {"dyntype": {"none": DYN_NONE, "integrator": DYN_INTEGRATOR, "filter": DYN_FILTER,
                             "filterexact": DYN_FILTEREXACT, "muscle": DYN_MUSCLE},
                 "gaintype": {"fixed": GAIN_FIXED, "affine": GAIN_AFFINE, "muscle": GAIN_MUSCLE},
                 "biastype": {"none": BIAS_NONE, "affine": BIAS_AFFINE, "muscle": BIAS_MUSCLE}}
        for key, table in names.items():
            if key in a:
                d[key] = table[a[key]]
        for key in ("dynprm", "gainprm", "biasprm"):
            if key in a:
                v = _floats(a[key])[:NPRM]
                d[key][:v.size] = v
    elif tag == "muscle":
        d["dyntype"], d["gaintype"], d["biastype"] = DYN_MUSCLE, GAIN_MUSCLE, BIAS_MUSCLE
        # muscle defaults replace untouched global defaults
        if d["dynprm"][0] == 1:
            d["dynprm"][0] = 0.01
        if d["dynprm"][1] == 0:
            d["dynprm"][1] = 0.04
        for i, (glob, val) in enumerate(((1, 0.75), (0, 1.05), (0, -1.0), (0, 200.0), (0, 0.5), (0, 1.6), (0, 1.5), (0, 1.3), (0, 1.2))):
            if d["gainprm"][i] == glob:
                d["gainprm"][i] = val
        if "timeconst" in a:
            d["dynprm"][0:2] = _floats(a["timeconst"], 2)
        if "tausmooth" in a:
            d["dynprm"][2] = float(a["tausmooth"])
        if "range" in a:
            d["gainprm"][0:2] = _floats(a["range"], 2)
        for i, key in ((2, "force"), (3, "scale"), (4, "lmin"), (5, "lmax"), (6, "vmax"), (7, "fpmax"), (8, "fvmax")):
            if key in a:
                d["gainprm"][i] = float(a[key])
        d["biasprm"][:9] = d["gainprm"][:9]
    elif tag == "position":
        d["dyntype"], d["gaintype"], d["biastype"] = DYN_NONE, GAIN_FIXED, BIAS_AFFINE
        if "kp" in a:
            d["gainprm"][0] = float(a["kp"])
        d["biasprm"][:3] = [0.0, -d["gainprm"][0], -float(a.get("kv", 0.0))]
    elif tag == "velocity":
        d["dyntype"], d["gaintype"], d["biastype"] = DYN_NONE, GAIN_FIXED, BIAS_AFFINE
        if "kv" in a:
            d["gainprm"][0] = float(a["kv"])
        d["biasprm"][:3] = [0.0, 0.0, -d["gainprm"][0]]
    elif tag == "motor":
        pass
    return d


class _Defaults:
    def __init__(self):
        self.cls = {"main": {"_act": _actuator_global_default()}}

    def load(self, elem, parent="main", top=True):
        name = elem.attrib.get("class", "main" if top else None)
        if name is None:
            raise MjcfError("nested <default> needs a class")
        if name not in self.cls:
            base = self.cls[parent]
            self.cls[name] = {k: (_copy_act(v) if k == "_act" else dict(v)) for k, v in base.items()}
        cur = self.cls[name]
        for c in elem:
            if c.tag == "default":
                continue
            if c.tag in _ACT_TAGS:
                cur["_act"] = _apply_actuator(c.tag, c.attrib, cur["_act"])
                # non-numeric attributes (tendon=..., class=...) are not defaults
            else:
                cur.setdefault(c.tag, {}).update(c.attrib)
        for c in elem:
            if c.tag == "default":
                self.load(c, name, top=False)

    def attrs(self, tag, elem, childclass):
        cname = elem.attrib.get("class", childclass or "main")
        if cname not in self.cls:
            raise MjcfError(f"unknown default class {cname}")
        out = dict(self.cls[cname].get(tag, {}))
        out.update(elem.attrib)
        return out

    def act(self, elem):
        cname = elem.attrib.get("class", "main")
        return self.cls[cname]["_act"]


# ---- the compiler ----------------------------------------------------------------------------------
class Model:
    """Bag of numpy arrays named like mjModel fields (+ `opt_*`, `stat_meaninertia`, `names_*`)."""

    def save(self, path):
        np.savez_compressed(path, **{k: v for k, v in self.__dict__.items()})

    @staticmethod
    def load(path):
        m = Model()
        with np.load(path, allow_pickle=False) as z:
            for k in z.files:
                v = z[k]
                m.__dict__[k] = v.item() if v.ndim == 0 and v.dtype.kind in "iuf" else v
        if "geom_dataid" not in m.__dict__:   # fixtures written before mesh colliders existed
            m.nmesh, m.nmeshvert = 0, 0
            m.mesh_vertadr = np.zeros(0, dtype=np.int32); m.mesh_vertnum = np.zeros(0, dtype=np.int32)
            m.mesh_vert = np.zeros((0, 3)); m.geom_dataid = np.full(int(m.ngeom), -1, dtype=np.int32)
        return m

    def name2id(self, kind, name):
        names = [str(s) for s in getattr(self, "names_" + kind)]
        return names.index(name)


class _Body:
    def __init__(self):
        self.name = ""
        self.parent = -1
        self.pos = np.zeros(3)
        self.quat = np.array([1.0, 0, 0, 0])
        self.inertial = None  # (pos, quat, mass, diag)
        self.joints, self.geoms, self.sites = [], [], []


def from_xml_path(path, **overrides) -> Model:
    """Compile an MJCF file.  `overrides` may set `timestep`, `iterations`, `ls_iterations`
    (the runtime edits of myosuite/mjx/myodm_v0.py:38-41)."""
    path = os.path.abspath(path)
    main_dir = os.path.dirname(path)
    root = _load_tree(path, main_dir)
    return _Compiler(root, main_dir).compile(overrides)


class _Compiler:
    def __init__(self, root, main_dir):
        self.root = root
        self.main_dir = main_dir
        self.comp = dict(angle="degree", eulerseq="xyz", inertiafromgeom="auto", balanceinertia="false",
                         boundmass="0", boundinertia="0", meshdir="", autolimits="true", coordinate="local",
                         settotalmass="-1")
        self.opt = dict(timestep=0.002, gravity=np.array([0, 0, -9.81]), iterations=100, ls_iterations=50,
                        tolerance=1e-8, ls_tolerance=0.01, impratio=1.0, cone=0, solver=2, integrator=0,
                        disableflags=0, noslip_iterations=0)
        self.defaults = _Defaults()
        self.meshes = {}
        self.mesh_verts = {}
        self.mesh_cache = {}
        self.bodies = []

    # -- attribute helpers
    def _angle(self, v):
        return v * (math.pi / 180.0) if self.comp["angle"] == "degree" else v

    def _orient(self, a):
        if "quat" in a:
            q = _floats(a["quat"], 4)
            return q / np.linalg.norm(q)
        if "euler" in a:
            e = self._angle(_floats(a["euler"], 3))
            q = np.array([1.0, 0, 0, 0])
            for ch, ang in zip(self.comp["eulerseq"], e):
                ax = np.zeros(3)
                ax["xyz".index(ch.lower())] = 1.0
                r = quat_from_axis_angle(ax, ang)
                q = quat_mul(q, r) if ch.islower() else quat_mul(r, q)
            return q
        if "axisangle" in a:
            v = _floats(a["axisangle"], 4)
            return quat_from_axis_angle(v[:3] / np.linalg.norm(v[:3]), self._angle(v[3]))
        if "zaxis" in a:
            return quat_z_to_vec(_floats(a["zaxis"], 3))
        if "xyaxes" in a:
            v = _floats(a["xyaxes"], 6)
            x = v[:3] / np.linalg.norm(v[:3])
            y = v[3:] - x * np.dot(x, v[3:])
            y /= np.linalg.norm(y)
            return mat_to_quat(np.stack([x, y, np.cross(x, y)], axis=1))
        return np.array([1.0, 0, 0, 0])

    # -- top level
    def compile(self, overrides):
        r = self.root
        for c in r.iter("compiler"):
            self.comp.update(c.attrib)
        if self.comp["coordinate"] != "local":
            raise MjcfError("coordinate=global")
        for o in r.iter("option"):
            a = o.attrib
            for k in ("timestep", "tolerance", "ls_tolerance", "impratio"):
                if k in a:
                    self.opt[k] = float(a[k])
            for k in ("iterations", "ls_iterations", "noslip_iterations"):
                if k in a:
                    self.opt[k] = int(a[k])
            if "gravity" in a:
                self.opt["gravity"] = _floats(a["gravity"], 3)
            if "cone" in a:
                self.opt["cone"] = {"pyramidal": 0, "elliptic": 1}[a["cone"]]
            if "solver" in a:
                self.opt["solver"] = {"PGS": 0, "CG": 1, "Newton": 2}[a["solver"]]
            if "integrator" in a and a["integrator"] != "Euler":
                raise MjcfError("integrator " + a["integrator"])
            for f in o.iter("flag"):
                bits = dict(constraint=0, equality=1, frictionloss=2, limit=3, contact=4, passive=5, gravity=6,
                            clampctrl=7, warmstart=8, filterparent=9, actuation=10, refsafe=11, sensor=12,
                            midphase=13, eulerdamp=14)
                for k, v in f.attrib.items():
                    if k in bits and v == "disable":
                        self.opt["disableflags"] |= 1 << bits[k]
        for k, v in overrides.items():
            if k not in self.opt:
                raise MjcfError("unknown override " + k)
            self.opt[k] = v
        if self.opt["cone"] != 0:
            raise MjcfError("elliptic cones")
        for d in r.findall("default"):
            self.defaults.load(d)
        for asset in r.findall("asset"):
            for me in asset.findall("mesh"):
                a = self.defaults.attrs("mesh", me, None)
                name = a.get("name") or os.path.splitext(os.path.basename(a["file"]))[0]
                self.meshes[name] = (a.get("file"), _floats(a.get("scale", "1 1 1"), 3))
                if "vertex" in a:   # inline vertex list: the mesh is its convex hull (as in MuJoCo)
                    self.mesh_verts[name] = _floats(a["vertex"]).reshape(-1, 3) * _floats(a.get("scale", "1 1 1"), 3)

        world = _Body()
        world.name = "world"
        self.bodies.append(world)
        for wb in r.findall("worldbody"):
            self._body_children(wb, 0, None)
        return self._finish()

    # -- kinematic tree (bodies are numbered depth-first; elements in document order inside a body)
    def _body_children(self, elem, bid, childclass):
        pending = []
        for c in elem:
            if c.tag == "body":
                pending.append(c)
            elif c.tag == "inertial":
                a = c.attrib
                pos = _floats(a["pos"], 3)
                mass = float(a["mass"])
                if "fullinertia" in a:
                    f = _floats(a["fullinertia"], 6)
                    I = np.array([[f[0], f[3], f[4]], [f[3], f[1], f[5]], [f[4], f[5], f[2]]])
                    diag, q = principal_axes(I)
                    q = quat_mul(self._orient(a), q) if ("quat" in a or "euler" in a) else q
                else:
                    diag = _floats(a["diaginertia"], 3) if "diaginertia" in a else np.zeros(3)
                    q = self._orient(a)
                self.bodies[bid].inertial = (pos, q, mass, diag)
            elif c.tag in ("joint", "freejoint"):
                a = self.defaults.attrs("joint", c, childclass) if c.tag == "joint" else dict(c.attrib, type="free")
                self.bodies[bid].joints.append(a)
            elif c.tag == "geom":
                self.bodies[bid].geoms.append(self.defaults.attrs("geom", c, childclass))
            elif c.tag == "site":
                self.bodies[bid].sites.append(self.defaults.attrs("site", c, childclass))
            elif c.tag in ("camera", "light"):
                pass
            else:
                raise MjcfError(f"<{c.tag}> inside body")
        # children are numbered after this body's own elements are known; recursion keeps DFS order
        for c in pending:
            b = _Body()
            b.name = c.attrib.get("name", f"body{len(self.bodies)}")
            b.parent = bid
            b.pos = _floats(c.attrib.get("pos", "0 0 0"), 3)
            b.quat = self._orient(c.attrib)
            if c.attrib.get("mocap", "false") == "true":
                raise MjcfError("mocap bodies")
            self.bodies.append(b)
            self._body_children(c, len(self.bodies) - 1, c.attrib.get("childclass", childclass))

    # -- geoms
    def _mesh_props(self, name):
        """(volume, centre of mass, principal inertia, principal-axes quaternion, hull vertices in the re-centred frame)"""
        if name not in self.mesh_cache:
            f, scale = self.meshes[name]
            tri = None
            if name in self.mesh_verts:
                tri = hull_triangles(self.mesh_verts[name])
            elif f is not None:
                p = os.path.normpath(os.path.join(self.main_dir, self.comp["meshdir"], f))
                if p.lower().endswith(".stl"):
                    tri = load_stl(p, scale)
            if tri is None:
                self.mesh_cache[name] = None  # .msh/.obj scene props: static, visual only
            else:
                vol, com, I = mesh_mass_properties(tri)
                diag, q = principal_axes(I)
                # collision geometry = the convex hull (MuJoCo: mjCMesh::MakeGraph), in the mesh frame MuJoCo re-centres on the
                # centre of mass and the principal axes
                hv = hull_vertices(tri.reshape(-1, 3))
                hv = (hv - com) @ quat_to_mat(q)
                self.mesh_cache[name] = (vol, com, diag, q, hv)
        return self.mesh_cache[name]

    def _geom(self, a, need_mass):
        t = GEOM_NAMES[a.get("type", "sphere")] if "mesh" not in a or "type" in a else GEOM_MESH
        if "mesh" in a and "type" not in a:
            t = GEOM_MESH
        size = np.zeros(3)
        sv = _floats(a.get("size", "0"))
        size[:min(3, sv.size)] = sv[:3]
        pos = _floats(a.get("pos", "0 0 0"), 3)
        quat = self._orient(a)
        if "fromto" in a:
            ft = _floats(a["fromto"], 6)
            vec = ft[0:3] - ft[3:6]
            pos = 0.5 * (ft[0:3] + ft[3:6])
            quat = quat_z_to_vec(vec)
            size[1] = 0.5 * np.linalg.norm(vec)
            if t == GEOM_BOX or t == GEOM_ELLIPSOID:
                size[2] = size[1]
                size[1] = size[0]
        vol, inertia_unit = 0.0, np.zeros(3)  # inertia per unit mass about geom frame
        r, h = size[0], size[1]
        if t == GEOM_SPHERE:
            vol = 4.0 / 3.0 * math.pi * r ** 3
            inertia_unit[:] = 0.4 * r * r
            rbound = r
        elif t == GEOM_CAPSULE:
            vc, vs = math.pi * r * r * 2 * h, 4.0 / 3.0 * math.pi * r ** 3
            vol = vc + vs
            mc, ms = vc / vol, vs / vol
            ixx = mc * (3 * r * r + 4 * h * h) / 12.0 + ms * (0.4 * r * r + h * h + 0.75 * h * r)
            inertia_unit[:] = [ixx, ixx, mc * r * r / 2 + ms * 0.4 * r * r]
            rbound = r + h
        elif t == GEOM_CYLINDER:
            vol = math.pi * r * r * 2 * h
            ixx = (3 * r * r + 4 * h * h) / 12.0
            inertia_unit[:] = [ixx, ixx, r * r / 2]
            rbound = math.sqrt(r * r + h * h)
        elif t == GEOM_ELLIPSOID:
            vol = 4.0 / 3.0 * math.pi * size[0] * size[1] * size[2]
            inertia_unit[:] = [(size[1] ** 2 + size[2] ** 2) / 5, (size[0] ** 2 + size[2] ** 2) / 5, (size[0] ** 2 + size[1] ** 2) / 5]
            rbound = size.max()
        elif t == GEOM_BOX:
            vol = 8 * size[0] * size[1] * size[2]
            inertia_unit[:] = [(size[1] ** 2 + size[2] ** 2) / 3, (size[0] ** 2 + size[2] ** 2) / 3, (size[0] ** 2 + size[1] ** 2) / 3]
            rbound = np.linalg.norm(size)
        elif t == GEOM_MESH:
            props = self._mesh_props(a["mesh"]) if (need_mass or a["mesh"] in self.mesh_cache or self.meshes[a["mesh"]][0].lower().endswith(".stl")) else None
            rbound = 0.0
            if props is not None:
                mvol, mcom, mdiag, mq, mhull = props
                rbound = float(np.linalg.norm(mhull, axis=1).max())   # (geom_size stays zero: collision reads the hull)
                # the mesh frame is re-centred on its centre of mass / principal axes
                pos = pos + quat_to_mat(quat) @ mcom
                quat = quat_mul(quat, mq)
                vol = mvol
                inertia_unit[:] = mdiag / max(mvol, MINVAL)
        else:  # plane, hfield
            rbound = 0.0
        if "mass" in a:
            mass = float(a["mass"])
        else:
            mass = float(a.get("density", 1000.0)) * vol
        g = dict(type=t, size=size, pos=pos, quat=quat / np.linalg.norm(quat), mass=mass, inertia=inertia_unit * mass,
                 rbound=rbound, name=a.get("name", ""), mesh=(a["mesh"] if t == GEOM_MESH and self._mesh_props(a["mesh"]) is not None else None),
                 contype=int(a.get("contype", 1)), conaffinity=int(a.get("conaffinity", 1)), condim=int(a.get("condim", 3)),
                 priority=int(a.get("priority", 0)), margin=float(a.get("margin", 0)), gap=float(a.get("gap", 0)),
                 solmix=float(a.get("solmix", 1)), group=int(a.get("group", 0)))
        fr = _floats(a.get("friction", "1 0.005 0.0001"))
        g["friction"] = np.array([1, 0.005, 0.0001])
        g["friction"][:fr.size] = fr
        sr = _floats(a.get("solref", "0.02 1"))
        g["solref"] = sr
        si = np.array([0.9, 0.95, 0.001, 0.5, 2.0])
        sv = _floats(a["solimp"]) if "solimp" in a else si
        si[:sv.size] = sv
        g["solimp"] = si
        return g

    # -- finish: flatten everything
    def _finish(self):
        m = Model()
        B = self.bodies
        nbody = len(B)
        auto = self.comp["autolimits"] == "true"

        def limited(a, rng_key="range"):
            v = a.get("limited", "auto")
            if v == "auto":
                if not auto:
                    return False
                return rng_key in a
            return v == "true"

        # bodies
        m.nbody = nbody
        m.names_body = np.array([b.name for b in B])
        m.body_parentid = np.array([max(b.parent, 0) for b in B], dtype=np.int32)
        m.body_pos = np.array([b.pos for b in B])
        m.body_quat = np.array([b.quat for b in B])

        # joints / dofs
        jnt = dict(type=[], qposadr=[], dofadr=[], bodyid=[], pos=[], axis=[], range=[], limited=[], stiffness=[],
                   margin=[], solref=[], solimp=[], qpos0=[], name=[])
        dof = dict(bodyid=[], jntid=[], parentid=[], damping=[], armature=[], frictionloss=[])
        qpos0, qpos_spring = [], []
        body_jntadr, body_jntnum, body_dofadr, body_dofnum = [], [], [], []
        last_dof_of_body = {}
        nq = nv = 0
        for bid, b in enumerate(B):
            body_jntadr.append(len(jnt["type"]) if b.joints else -1)
            body_jntnum.append(len(b.joints))
            body_dofadr.append(nv if b.joints else -1)
            nd0 = nv
            # parent dof = last dof of the nearest ancestor that has dofs
            p = b.parent
            pd = -1
            while p > 0:
                if p in last_dof_of_body:
                    pd = last_dof_of_body[p]
                    break
                p = B[p].parent
            for a in b.joints:
                tname = a.get("type", "hinge")
                t = {"free": JNT_FREE, "ball": JNT_BALL, "slide": JNT_SLIDE, "hinge": JNT_HINGE}[tname]
                if t == JNT_BALL:
                    raise MjcfError("ball joints")
                jid = len(jnt["type"])
                jnt["type"].append(t)
                jnt["name"].append(a.get("name", f"joint{jid}"))
                jnt["qposadr"].append(nq)
                jnt["dofadr"].append(nv)
                jnt["bodyid"].append(bid)
                jnt["pos"].append(_floats(a.get("pos", "0 0 0"), 3))
                ax = _floats(a.get("axis", "0 0 1"), 3)
                jnt["axis"].append(ax / np.linalg.norm(ax))
                rng = _floats(a.get("range", "0 0"), 2)
                lim = limited(a) and t != JNT_FREE
                if t == JNT_HINGE:
                    rng = self._angle(rng)
                jnt["range"].append(rng)
                jnt["limited"].append(lim)
                jnt["stiffness"].append(float(a.get("stiffness", 0)))
                jnt["margin"].append(float(a.get("margin", 0)))
                jnt["solref"].append(_floats(a.get("solreflimit", "0.02 1"), 2))
                si = np.array([0.9, 0.95, 0.001, 0.5, 2.0])
                if "solimplimit" in a:
                    sv = _floats(a["solimplimit"])
                    si[:sv.size] = sv
                jnt["solimp"].append(si)
                if t == JNT_FREE:
                    q0 = np.concatenate([b.pos, b.quat])
                    ndof = 6
                else:
                    ref = float(a.get("ref", 0))
                    q0 = np.array([self._angle(ref) if t == JNT_HINGE else ref])
                    ndof = 1
                qpos0.append(q0)
                sp = float(a.get("springref", 0))
                qpos_spring.append(q0 if t == JNT_FREE else np.array([self._angle(sp) if t == JNT_HINGE else sp]))
                for k in range(ndof):
                    dof["bodyid"].append(bid)
                    dof["jntid"].append(jid)
                    dof["parentid"].append(pd)
                    pd = nv
                    dof["damping"].append(float(a.get("damping", 0)))
                    dof["armature"].append(float(a.get("armature", 0)))
                    dof["frictionloss"].append(float(a.get("frictionloss", 0)))
                    nv += 1
                nq += q0.size
            if nv > nd0:
                last_dof_of_body[bid] = nv - 1
            body_dofnum.append(nv - nd0)
        if any(f != 0 for f in dof["frictionloss"]):
            raise MjcfError("joint frictionloss")
        m.nq, m.nv, m.njnt = nq, nv, len(jnt["type"])
        m.names_jnt = np.array(jnt["name"])
        m.jnt_type = np.array(jnt["type"], dtype=np.int32)
        m.jnt_qposadr = np.array(jnt["qposadr"], dtype=np.int32)
        m.jnt_dofadr = np.array(jnt["dofadr"], dtype=np.int32)
        m.jnt_bodyid = np.array(jnt["bodyid"], dtype=np.int32)
        m.jnt_pos = np.array(jnt["pos"]).reshape(-1, 3)
        m.jnt_axis = np.array(jnt["axis"]).reshape(-1, 3)
        m.jnt_range = np.array(jnt["range"]).reshape(-1, 2)
        m.jnt_limited = np.array(jnt["limited"], dtype=np.int32)
        m.jnt_stiffness = np.array(jnt["stiffness"])
        m.jnt_margin = np.array(jnt["margin"])
        m.jnt_solref = np.array(jnt["solref"]).reshape(-1, 2)
        m.jnt_solimp = np.array(jnt["solimp"]).reshape(-1, 5)
        m.qpos0 = np.concatenate(qpos0) if qpos0 else np.zeros(0)
        m.qpos_spring = np.concatenate(qpos_spring) if qpos_spring else np.zeros(0)
        m.body_jntadr = np.array(body_jntadr, dtype=np.int32)
        m.body_jntnum = np.array(body_jntnum, dtype=np.int32)
        m.body_dofadr = np.array(body_dofadr, dtype=np.int32)
        m.body_dofnum = np.array(body_dofnum, dtype=np.int32)
        m.dof_bodyid = np.array(dof["bodyid"], dtype=np.int32)
        m.dof_jntid = np.array(dof["jntid"], dtype=np.int32)
        m.dof_parentid = np.array(dof["parentid"], dtype=np.int32)
        m.dof_damping = np.array(dof["damping"])
        m.dof_armature = np.array(dof["armature"])
        # sparse-M addressing: row i holds (i,i),(i,parent(i)),...
        madr, nM = [], 0
        for i in range(nv):
            madr.append(nM)
            j = i
            while j >= 0:
                nM += 1
                j = dof["parentid"][j]
        m.dof_Madr = np.array(madr, dtype=np.int32)
        m.nM = nM
        # weld / root ids
        weld, rootid = [0] * nbody, [0] * nbody
        for bid in range(1, nbody):
            p = B[bid].parent
            weld[bid] = bid if B[bid].joints else weld[p]
            rootid[bid] = bid if p == 0 else rootid[p]
        m.body_weldid = np.array(weld, dtype=np.int32)
        m.body_rootid = np.array(rootid, dtype=np.int32)

        # geoms (+ inertia inference), sites
        G, S = [], []
        body_geomadr, body_geomnum = [], []
        ipos, iquat, mass, inertia = [], [], [], []
        infer_mode = self.comp["inertiafromgeom"]
        for bid, b in enumerate(B):
            infer = (infer_mode == "true") or (infer_mode == "auto" and b.inertial is None)
            body_geomadr.append(len(G) if b.geoms else -1)
            body_geomnum.append(len(b.geoms))
            gl = []
            for a in b.geoms:
                g = self._geom(a, need_mass=infer and bid > 0)
                g["bodyid"] = bid
                gl.append(g)
                G.append(g)
            for a in b.sites:
                sz = np.zeros(3)
                sv = _floats(a.get("size", "0.005"))
                sz[:sv.size] = sv
                S.append(dict(bodyid=bid, pos=_floats(a.get("pos", "0 0 0"), 3), quat=self._orient(a), name=a.get("name", ""),
                              size=sz, type=GEOM_NAMES[a.get("type", "sphere")]))
            if bid == 0:
                ipos.append(np.zeros(3)); iquat.append(np.array([1.0, 0, 0, 0])); mass.append(0.0); inertia.append(np.zeros(3))
                continue
            if infer and gl:
                mt = sum(g["mass"] for g in gl)
                if mt > MINVAL:
                    com = sum(g["mass"] * g["pos"] for g in gl) / mt
                    I = np.zeros((3, 3))
                    for g in gl:
                        R = quat_to_mat(g["quat"])
                        d = g["pos"] - com
                        I += R @ np.diag(g["inertia"]) @ R.T + g["mass"] * (np.dot(d, d) * np.eye(3) - np.outer(d, d))
                    diag, q = principal_axes(I)
                else:
                    com, diag, q = np.zeros(3), np.zeros(3), np.array([1.0, 0, 0, 0])
                bi = (com, q, mt, diag)
            elif b.inertial is not None:
                bi = b.inertial
            else:
                bi = (np.zeros(3), np.array([1.0, 0, 0, 0]), 0.0, np.zeros(3))
            p, q, ms, dg = bi
            dg = np.array(dg, dtype=np.float64)
            if self.comp["balanceinertia"] == "true":
                if dg[0] + dg[1] < dg[2] or dg[0] + dg[2] < dg[1] or dg[1] + dg[2] < dg[0]:
                    dg[:] = dg.mean()
            ms = max(ms, float(self.comp["boundmass"]))
            dg = np.maximum(dg, float(self.comp["boundinertia"]))
            ipos.append(np.asarray(p, dtype=np.float64)); iquat.append(q); mass.append(ms); inertia.append(dg)
        m.body_ipos = np.array(ipos)
        m.body_iquat = np.array(iquat)
        m.body_mass = np.array(mass)
        m.body_inertia = np.array(inertia)
        for bid in range(1, nbody):
            if m.body_dofnum[bid] > 0 and (m.body_mass[bid] < MINVAL or m.body_inertia[bid].min() < MINVAL):
                # MuJoCo only rejects this when no descendant carries mass; the MyoSuite models never hit it
                sub = [c for c in range(bid, nbody) if self._is_descendant(c, bid)]
                if sum(m.body_mass[c] for c in sub) < MINVAL:
                    raise MjcfError(f"moving body {B[bid].name} has no mass")
        sub = m.body_mass.copy()
        for bid in range(nbody - 1, 0, -1):
            sub[m.body_parentid[bid]] += sub[bid]
        m.body_subtreemass = sub
        m.body_geomadr = np.array(body_geomadr, dtype=np.int32)
        m.body_geomnum = np.array(body_geomnum, dtype=np.int32)

        m.ngeom = len(G)
        m.names_geom = np.array([g["name"] for g in G])
        for key, dt in (("type", np.int32), ("bodyid", np.int32), ("contype", np.int32), ("conaffinity", np.int32),
                        ("condim", np.int32), ("priority", np.int32), ("group", np.int32), ("size", float), ("pos", float), ("quat", float),
                        ("rbound", float), ("margin", float), ("gap", float), ("solmix", float), ("friction", float),
                        ("solref", float), ("solimp", float)):
            arr = np.array([g[key] for g in G], dtype=dt)
            if arr.ndim == 1 and key in ("size", "pos", "quat", "friction", "solref", "solimp"):
                arr = arr.reshape(0, {"size": 3, "pos": 3, "quat": 4, "friction": 3, "solref": 2, "solimp": 5}[key])
            setattr(m, "geom_" + key, arr)
        # convex hulls of the mesh geoms (mjModel: mesh_vertadr / mesh_vertnum / mesh_vert, geom_dataid)
        mesh_ids, adr, num, verts = {}, [], [], []
        dataid = np.full(len(G), -1, dtype=np.int32)
        for gi, g in enumerate(G):
            if g.get("mesh") is None:
                continue
            if g["mesh"] not in mesh_ids:
                hv = self._mesh_props(g["mesh"])[4]
                mesh_ids[g["mesh"]] = len(adr)
                adr.append(sum(num)); num.append(len(hv)); verts.append(hv)
            dataid[gi] = mesh_ids[g["mesh"]]
        m.nmesh, m.nmeshvert = len(adr), int(sum(num))
        m.mesh_vertadr = np.array(adr, dtype=np.int32)
        m.mesh_vertnum = np.array(num, dtype=np.int32)
        m.mesh_vert = np.concatenate(verts).reshape(-1, 3) if verts else np.zeros((0, 3))
        m.geom_dataid = dataid
        m.nsite = len(S)
        m.names_site = np.array([s["name"] for s in S])
        m.site_bodyid = np.array([s["bodyid"] for s in S], dtype=np.int32)
        m.site_pos = np.array([s["pos"] for s in S]).reshape(-1, 3)
        m.site_quat = np.array([s["quat"] for s in S]).reshape(-1, 4)
        m.site_size = np.array([s["size"] for s in S]).reshape(-1, 3)
        m.site_type = np.array([s["type"] for s in S], dtype=np.int32)

        self._tendons(m, limited)
        self._actuators(m)
        self._contacts(m)
        self._equalities(m)
        self._keys_sensors(m)

        for k, v in self.opt.items():
            setattr(m, "opt_" + k, v)
        setconst.set_const(m)
        return m

    def _is_descendant(self, c, b):
        while c > 0:
            if c == b:
                return True
            c = self.bodies[c].parent
        return False

    def _tendons(self, m, limited):
        wt, wo, wp = [], [], []
        T = []
        for sec in self.root.findall("tendon"):
            for te in sec:
                if te.tag != "spatial":
                    raise MjcfError(f"<tendon><{te.tag}>")
                a = self.defaults.attrs("tendon", te, None)
                adr = len(wt)
                for w in te:
                    if w.tag == "site":
                        wt.append(WRAP_SITE); wo.append(m.name2id("site", w.attrib["site"])); wp.append(0.0)
                    elif w.tag == "geom":
                        gid = m.name2id("geom", w.attrib["geom"])
                        gt = m.geom_type[gid]
                        if gt not in (GEOM_SPHERE, GEOM_CYLINDER):
                            raise MjcfError("tendon wraps only spheres and cylinders")
                        wt.append(WRAP_SPHERE if gt == GEOM_SPHERE else WRAP_CYLINDER)
                        wo.append(gid)
                        wp.append(float(m.name2id("site", w.attrib["sidesite"])) if "sidesite" in w.attrib else -1.0)
                    elif w.tag == "pulley":
                        wt.append(WRAP_PULLEY); wo.append(-1); wp.append(float(w.attrib["divisor"]))
                    else:
                        raise MjcfError(f"tendon <{w.tag}>")
                num = len(wt) - adr
                types = wt[adr:]
                if num < 2 or types[0] != WRAP_SITE or types[-1] != WRAP_SITE:
                    raise MjcfError("spatial tendon must start and end with a site")
                sl = _floats(a.get("springlength", "-1"))
                T.append(dict(name=a.get("name", ""), adr=adr, num=num, limited=limited(a), range=_floats(a.get("range", "0 0"), 2),
                              stiffness=float(a.get("stiffness", 0)), damping=float(a.get("damping", 0)),
                              frictionloss=float(a.get("frictionloss", 0)), margin=float(a.get("margin", 0)),
                              springlength=np.array([sl[0], sl[-1]]),
                              solref=_floats(a.get("solreflimit", "0.02 1"), 2),
                              solimp=_pad(_floats(a["solimplimit"]) if "solimplimit" in a else np.zeros(0), [0.9, 0.95, 0.001, 0.5, 2.0])))
        m.ntendon = len(T)
        m.nwrap = len(wt)
        m.names_tendon = np.array([t["name"] for t in T])
        m.tendon_adr = np.array([t["adr"] for t in T], dtype=np.int32)
        m.tendon_num = np.array([t["num"] for t in T], dtype=np.int32)
        m.tendon_limited = np.array([t["limited"] for t in T], dtype=np.int32)
        m.tendon_range = np.array([t["range"] for t in T]).reshape(-1, 2)
        m.tendon_stiffness = np.array([t["stiffness"] for t in T])
        m.tendon_damping = np.array([t["damping"] for t in T])
        m.tendon_margin = np.array([t["margin"] for t in T])
        m.tendon_lengthspring = np.array([t["springlength"] for t in T]).reshape(-1, 2)
        m.tendon_solref_lim = np.array([t["solref"] for t in T]).reshape(-1, 2)
        m.tendon_solimp_lim = np.array([t["solimp"] for t in T]).reshape(-1, 5)
        if any(t["frictionloss"] != 0 for t in T):
            raise MjcfError("tendon frictionloss")
        m.wrap_type = np.array(wt, dtype=np.int32)
        m.wrap_objid = np.array(wo, dtype=np.int32)
        m.wrap_prm = np.array(wp, dtype=np.float64)

    def _actuators(self, m):
        A = []
        for sec in self.root.findall("actuator"):
            for e in sec:
                if e.tag not in _ACT_TAGS:
                    raise MjcfError(f"<actuator><{e.tag}>")
                d = _apply_actuator(e.tag, e.attrib, self.defaults.act(e))
                d["name"] = e.attrib.get("name", "")
                if "tendon" in e.attrib:
                    d["trntype"], d["trnid"] = TRN_TENDON, m.name2id("tendon", e.attrib["tendon"])
                elif "joint" in e.attrib:
                    d["trntype"], d["trnid"] = TRN_JOINT, m.name2id("jnt", e.attrib["joint"])
                else:
                    raise MjcfError("actuator transmission")
                for key, rk in (("ctrllimited", "ctrlrange"), ("forcelimited", "forcerange"), ("actlimited", "actrange")):
                    if d[key] is None:
                        d[key] = bool(d[rk][0] != 0 or d[rk][1] != 0)
                A.append(d)
        m.nu = len(A)
        m.names_actuator = np.array([a["name"] for a in A])
        m.actuator_trntype = np.array([a["trntype"] for a in A], dtype=np.int32)
        m.actuator_trnid = np.array([a["trnid"] for a in A], dtype=np.int32)
        m.actuator_dyntype = np.array([a["dyntype"] for a in A], dtype=np.int32)
        m.actuator_gaintype = np.array([a["gaintype"] for a in A], dtype=np.int32)
        m.actuator_biastype = np.array([a["biastype"] for a in A], dtype=np.int32)
        m.actuator_dynprm = np.array([a["dynprm"] for a in A]).reshape(-1, NPRM)
        m.actuator_gainprm = np.array([a["gainprm"] for a in A]).reshape(-1, NPRM)
        m.actuator_biasprm = np.array([a["biasprm"] for a in A]).reshape(-1, NPRM)
        m.actuator_ctrllimited = np.array([a["ctrllimited"] for a in A], dtype=np.int32)
        m.actuator_ctrlrange = np.array([a["ctrlrange"] for a in A]).reshape(-1, 2)
        m.actuator_forcelimited = np.array([a["forcelimited"] for a in A], dtype=np.int32)
        m.actuator_forcerange = np.array([a["forcerange"] for a in A]).reshape(-1, 2)
        m.actuator_actlimited = np.array([a["actlimited"] for a in A], dtype=np.int32)
        m.actuator_actrange = np.array([a["actrange"] for a in A]).reshape(-1, 2)
        m.actuator_lengthrange = np.array([a["lengthrange"] for a in A]).reshape(-1, 2)
        m.actuator_gear = np.array([a["gear"][0] for a in A])
        # activation states: one per actuator with dynamics, in actuator order
        actadr, na = [], 0
        for a in A:
            if a["dyntype"] != DYN_NONE:
                actadr.append(na); na += 1
            else:
                actadr.append(-1)
        m.na = na
        m.actuator_actadr = np.array(actadr, dtype=np.int32)

    def _contacts(self, m):
        pairs, excl = [], []
        for sec in self.root.findall("contact"):
            for e in sec:
                if e.tag == "exclude":
                    b1, b2 = m.name2id("body", e.attrib["body1"]), m.name2id("body", e.attrib["body2"])
                    excl.append((min(b1, b2) << 16) + max(b1, b2))
                elif e.tag == "pair":
                    a = self.defaults.attrs("pair", e, None)
                    g1, g2 = m.name2id("geom", a["geom1"]), m.name2id("geom", a["geom2"])
                    pairs.append((g1, g2, a))
        m.exclude_signature = np.array(excl, dtype=np.int32)
        m.npair = len(pairs)
        m.pair_geom1 = np.array([p[0] for p in pairs], dtype=np.int32)
        m.pair_geom2 = np.array([p[1] for p in pairs], dtype=np.int32)
        dim, fr, sr, si, mg, gp = [], [], [], [], [], []
        for g1, g2, a in pairs:
            # unspecified pair parameters are mixed from the two geoms (same rule as dynamic pairs)
            mix = _mix_geoms(m, g1, g2)
            dim.append(int(a["condim"]) if "condim" in a else mix["condim"])
            f = mix["friction"].copy()
            if "friction" in a:
                fv = _floats(a["friction"])
                f[:fv.size] = fv
            fr.append(f)
            sr.append(_floats(a["solref"], 2) if "solref" in a else mix["solref"])
            si.append(_pad(_floats(a["solimp"]), mix["solimp"]) if "solimp" in a else mix["solimp"])
            mg.append(float(a["margin"]) if "margin" in a else mix["margin"])
            gp.append(float(a["gap"]) if "gap" in a else mix["gap"])
        m.pair_dim = np.array(dim, dtype=np.int32)
        m.pair_friction = np.array(fr).reshape(-1, 5)
        m.pair_solref = np.array(sr).reshape(-1, 2)
        m.pair_solimp = np.array(si).reshape(-1, 5)
        m.pair_margin = np.array(mg)
        m.pair_gap = np.array(gp)

    def _equalities(self, m):
        E = []
        for sec in self.root.findall("equality"):
            for e in sec:
                if e.tag != "joint":
                    raise MjcfError(f"<equality><{e.tag}>")
                a = self.defaults.attrs("equality", e, None)
                j1 = m.name2id("jnt", a["joint1"])
                j2 = m.name2id("jnt", a["joint2"]) if "joint2" in a else -1
                pc = _pad(_floats(a.get("polycoef", "0 1 0 0 0")), [0, 1, 0, 0, 0])
                E.append(dict(obj1=j1, obj2=j2, data=pc, active=a.get("active", "true") == "true",
                              solref=_floats(a.get("solref", "0.02 1"), 2),
                              solimp=_pad(_floats(a["solimp"]) if "solimp" in a else np.zeros(0), [0.9, 0.95, 0.001, 0.5, 2.0])))
        m.neq = len(E)
        m.eq_type = np.full(len(E), EQ_JOINT, dtype=np.int32)
        m.eq_obj1id = np.array([e["obj1"] for e in E], dtype=np.int32)
        m.eq_obj2id = np.array([e["obj2"] for e in E], dtype=np.int32)
        m.eq_active0 = np.array([e["active"] for e in E], dtype=np.int32)
        m.eq_data = np.array([e["data"] for e in E]).reshape(-1, 5)
        m.eq_solref = np.array([e["solref"] for e in E]).reshape(-1, 2)
        m.eq_solimp = np.array([e["solimp"] for e in E]).reshape(-1, 5)

    def _keys_sensors(self, m):
        kq, kv, ka = [], [], []
        for sec in self.root.findall("keyframe"):
            for k in sec.findall("key"):
                kq.append(_floats(k.attrib["qpos"], m.nq) if "qpos" in k.attrib else m.qpos0.copy())
                kv.append(_floats(k.attrib["qvel"], m.nv) if "qvel" in k.attrib else np.zeros(m.nv))
                ka.append(_floats(k.attrib["act"], m.na) if "act" in k.attrib else np.zeros(m.na))
        m.nkey = len(kq)
        m.key_qpos = np.array(kq, dtype=np.float64).reshape(len(kq), m.nq)
        m.key_qvel = np.array(kv, dtype=np.float64).reshape(len(kq), m.nv)
        m.key_act = np.array(ka, dtype=np.float64).reshape(len(kq), m.na)
        touch = []
        for sec in self.root.findall("sensor"):
            for s in sec:
                if s.tag != "touch":
                    raise MjcfError(f"<sensor><{s.tag}>")
                touch.append(m.name2id("site", s.attrib["site"]))
        m.nsensor = len(touch)
        m.sensor_objid = np.array(touch, dtype=np.int32)


def _pad(v, default):
    out = np.array(default, dtype=np.float64)
    out[:v.size] = v
    return out


def _mix_geoms(m, g1, g2):
    """Contact parameters of a dynamically generated pair (MuJoCo's mj_contactParam rule)."""
    p1, p2 = m.geom_priority[g1], m.geom_priority[g2]
    out = {}
    out["margin"] = max(m.geom_margin[g1], m.geom_margin[g2])
    out["gap"] = max(m.geom_gap[g1], m.geom_gap[g2])
    if p1 != p2:
        g = g1 if p1 > p2 else g2
        out["condim"] = int(m.geom_condim[g])
        f = m.geom_friction[g]
        out["solref"] = m.geom_solref[g].copy()
        out["solimp"] = m.geom_solimp[g].copy()
    else:
        out["condim"] = int(max(m.geom_condim[g1], m.geom_condim[g2]))
        f = np.maximum(m.geom_friction[g1], m.geom_friction[g2])
        s1, s2 = m.geom_solmix[g1], m.geom_solmix[g2]
        if s1 >= MINVAL and s2 >= MINVAL:
            mix = s1 / (s1 + s2)
        elif s1 < MINVAL and s2 < MINVAL:
            mix = 0.5
        elif s1 < MINVAL:
            mix = 0.0
        else:
            mix = 1.0
        r1, r2 = m.geom_solref[g1], m.geom_solref[g2]
        if r1[0] > 0 and r2[0] > 0:
            out["solref"] = mix * r1 + (1 - mix) * r2
        else:
            out["solref"] = np.minimum(r1, r2)
        out["solimp"] = mix * m.geom_solimp[g1] + (1 - mix) * m.geom_solimp[g2]
    out["friction"] = np.array([f[0], f[0], f[1], f[2], f[2]])
    return out
