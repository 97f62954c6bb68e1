"""Mesh-convex collision (first part): the convex hull of a mesh geom as a collider -- support function over
the hull's vertices inside the same MPR narrow phase as the other convex shapes.

CPU: the oracle on a cube given as a MESH (eight vertices) against the same cube as a BOX primitive: both go through MPR with
support functions that agree on a cube, so the contacts must coincide.  Compiler: hull, mass, inertia of the inline mesh.
GPU (marked): the CUDA path against the oracle on an irregular polytope resting on capsules, an ellipsoid and the plane."""
import os
import tempfile

import numpy as np
import pytest

CUBE = "-0.1 -0.07 -0.05  0.1 -0.07 -0.05  -0.1 0.07 -0.05  0.1 0.07 -0.05  -0.1 -0.07 0.05  0.1 -0.07 0.05  -0.1 0.07 0.05  0.1 0.07 0.05"
SCENE = """<mujoco><option timestep="0.002"/>
<asset><mesh name="blk" vertex="{verts}"/></asset>
<worldbody>
  <geom name="floor" type="plane" size="1 1 .1"/>
  <geom name="egg" type="ellipsoid" size="0.08 0.05 0.04" pos="0.05 0.02 0.2"/>
  <geom name="bar" type="capsule" size="0.03 0.15" pos="-0.08 0 0.22" euler="0 90 20"/>
  <geom name="ball" type="sphere" size="0.05" pos="0 -0.12 0.25"/>
  <body name="obj" pos="0 0 0.3"><freejoint/>{geom}</body>
</worldbody></mujoco>"""


def compile_scene(geom, verts=CUBE):
    from myosuite_mjx_b200 import mjcf
    d = tempfile.mkdtemp()
    p = os.path.join(d, "scene.xml")
    open(p, "w").write(SCENE.format(verts=verts, geom=geom))
    return mjcf.from_xml_path(p)


def poses(n, seed):
    rng = np.random.default_rng(seed)
    q = np.zeros((n, 7))
    q[:, :3] = np.array([0.0, 0.0, 0.27]) + rng.uniform(-0.06, 0.06, (n, 3))
    quat = rng.normal(size=(n, 4))
    q[:, 3:] = quat / np.linalg.norm(quat, axis=1, keepdims=True)
    return q


def test_compiler_builds_the_hull_and_mass_of_an_inline_mesh():
    m = compile_scene('<geom name="blk" type="mesh" mesh="blk" density="500"/>')
    assert m.nmesh == 1 and m.nmeshvert == 8 and m.geom_dataid[-1] == 0 and m.geom_type[-1] == 7
    assert abs(m.body_mass[1] - 500 * 0.2 * 0.14 * 0.1) < 1e-9
    I = np.sort(m.body_inertia[1])
    ref = np.sort(m.body_mass[1] / 12 * np.array([0.14 ** 2 + 0.1 ** 2, 0.2 ** 2 + 0.1 ** 2, 0.2 ** 2 + 0.14 ** 2]))
    assert np.allclose(I, ref, rtol=1e-9)
    assert abs(m.geom_rbound[-1] - np.linalg.norm([0.1, 0.07, 0.05])) < 1e-12
    # hull vertices in the re-centred principal frame: same set of pairwise distances as the cube's corners
    v = m.mesh_vert
    d = np.sort(np.linalg.norm(v[:, None] - v[None], axis=2).ravel())
    c = np.array([float(x) for x in CUBE.split()]).reshape(8, 3)
    assert np.allclose(d, np.sort(np.linalg.norm(c[:, None] - c[None], axis=2).ravel()), atol=1e-12)


def test_oracle_mesh_cube_collides_like_the_box_primitive():
    import oracle
    mb = compile_scene('<geom name="blk" type="box" size="0.1 0.07 0.05" density="500"/>')
    mm = compile_scene('<geom name="blk" type="mesh" mesh="blk" density="500"/>')
    ob, om = oracle.Oracle(mb), oracle.Oracle(mm)
    # the mesh geom's frame is the principal-axes frame: the same world pose of the BODY puts the same solid in the same place
    # only if the body frame maps to the same geom pose; both bodies carry the geom at the body origin up to that rotation
    hits = 0
    for q in poses(200, 5):
        for o in (ob, om):
            o.reset(); o.set_state(q, np.zeros(6), np.zeros(0), np.zeros(0)); o.forward()
        cb = [(tuple(g), d, p) for g, d, p in zip(np.array(ob.field("contact_geom")).reshape(-1, 2)[:ob.ncon], np.array(ob.field("contact_dist"))[:ob.ncon], np.array(ob.field("contact_pos")).reshape(-1, 3)[:ob.ncon]) if g[0] != 0]
        cm = [(tuple(g), d, p) for g, d, p in zip(np.array(om.field("contact_geom")).reshape(-1, 2)[:om.ncon], np.array(om.field("contact_dist"))[:om.ncon], np.array(om.field("contact_pos")).reshape(-1, 3)[:om.ncon]) if g[0] != 0]
        assert [c[0] for c in cb] == [c[0] for c in cm], (cb, cm)   # same convex pairs in the same order
        for (g, d0, p0), (_, d1, p1) in zip(cb, cm):
            assert abs(d0 - d1) < 1e-6 and np.abs(p0 - p1).max() < 1e-5, (g, d0, d1, p0, p1)
            hits += 1
        # plane: the box gives up to four corner contacts, the hull one (the deepest): same deepest distance
        pb = [d for g, d in zip(np.array(ob.field("contact_geom")).reshape(-1, 2)[:ob.ncon], np.array(ob.field("contact_dist"))[:ob.ncon]) if g[0] == 0]
        pm = [d for g, d in zip(np.array(om.field("contact_geom")).reshape(-1, 2)[:om.ncon], np.array(om.field("contact_dist"))[:om.ncon]) if g[0] == 0]
        assert (len(pb) > 0) == (len(pm) > 0)
        if pb:
            assert len(pm) == 1 and abs(min(pb) - pm[0]) < 1e-9
    assert hits > 50


POLY = " ".join("%.4f" % x for x in (np.random.default_rng(3).normal(size=(16, 3)) * np.array([0.09, 0.07, 0.05])).ravel())


@pytest.mark.gpu
def test_gpu_mesh_hull_contacts_match_the_oracle():
    """An irregular 16-point polytope (its hull) on a free body among an ellipsoid, a capsule, a sphere and the plane: contact
    count, geom pairs and order exact; distance / position to the convex narrow phase's tolerance (float32 poses: 2e-3 of the
    scene scale); constrained acceleration as in tests/test_parity_gpu.py."""
    import torch
    import oracle
    from myosuite_mjx_b200 import mjx
    m = compile_scene('<geom name="blk" type="mesh" mesh="blk" density="500"/>', verts=POLY)
    assert m.nmeshvert >= 8
    nw = 96
    q = poses(nw, 9)
    model = mjx.put_model(m)
    d = mjx.put_data(model, dict(qpos=q), nworld=nw, fields="all")
    mjx.forward(model, d)
    torch.cuda.synchronize()
    o = oracle.Oracle(m)
    ncon_total, mesh_pairs = 0, 0
    gid = len(m.geom_type) - 1
    for w in range(nw):
        o.reset(); o.set_state(q[w], np.zeros(6), np.zeros(0), np.zeros(0)); o.forward()
        n = o.ncon
        assert int(d.ncon[w, 0]) == n, f"world {w}: ncon {int(d.ncon[w, 0])} vs oracle {n}"
        g_ref = np.array(o.field("contact_geom")).reshape(-1, 2)[:n]
        assert np.array_equal(d.contact_geom[w].cpu().numpy().reshape(-1, 2)[:n], g_ref)
        if n:
            assert np.abs(d.contact_dist[w, :n].double().cpu().numpy() - np.array(o.field("contact_dist"))[:n]).max() < 2e-3 * 0.3
            assert np.abs(d.contact_pos[w].double().cpu().numpy().reshape(-1, 3)[:n] - np.array(o.field("contact_pos")).reshape(-1, 3)[:n]).max() < 2e-3 * 0.3
            sq = max(np.abs(o.qacc).max(), 1e-9)
            assert np.abs(d.qacc[w].double().cpu().numpy() - o.qacc).max() < 5e-3 * sq + 1e-3
        ncon_total += n
        mesh_pairs += int((g_ref == gid).any(axis=1).sum())
    assert mesh_pairs > 40 and ncon_total >= mesh_pairs
    # and a short rollout stays finite and unflagged
    mjx.step(model, d, 50)
    torch.cuda.synchronize()
    assert bool(torch.isfinite(d.qpos).all()) and int((d.status != 0).sum()) == 0
