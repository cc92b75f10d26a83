#!/usr/bin/env python
"""Offline model compiler: reference MJCF + STL assets -> hmj_model_t blob.

Reads (never copies) the reference's data files

  * simulation/assets/world.xml (+ <include> humanoid.xml)      -- kinematic tree, defaults, excludes
  * simulation/assets/humanoid_urdf/*.stl                        -- link meshes
  * mujoco_mpc/mjpc/tasks/humanoid_cap/stand/task.xml            -- cost terms / planner customs

and restates what MuJoCo's XML compiler (3.1.4; not in the reference tree, see
SURVEY.md F1) would put into mjModel for this one robot: class defaults, body /
joint / dof tables, inertial frames, mesh re-centring to the CoM / principal frame with
float32 vertices, convex hulls (Qhull "Qt" via scipy), bounding radii, the candidate
collision-pair list in contact order, qpos0, subtree masses, invweight0, meaninertia.

Output: humanoid-mujoco_b200/assets/humanoid_model.bin  (sizeof(hmj_model_t) bytes)
        humanoid-mujoco_b200/assets/humanoid_model.json (human-diffable dump)

The blob is committed (it has to travel to the GPU box, where /root/reference does
not exist); this script is the committed generator.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import xml.etree.ElementTree as ET

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from model_layout import DEFINES, MODEL_DTYPE  # noqa: E402

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT_REF = "/root/reference"
MJMINVAL = 1e-15


# ----------------------------------------------------------------------------- math
def quat_normalize(q):
    q = np.asarray(q, dtype=np.float64)
    return q / np.linalg.norm(q)


def quat_mul(a, b):
    return np.array([
        a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3],
        a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
        a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1],
        a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0],
    ])


def quat2mat(q):
    w, x, y, z = q
    return np.array([
        [w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y)],
        [2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x)],
        [2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z],
    ])


def mat2quat(R):
    # standard robust conversion, w >= 0 convention
    tr = np.trace(R)
    if tr > 0:
        s = np.sqrt(tr + 1.0) * 2
        q = [0.25 * s, (R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s]
    elif R[0, 0] > R[1, 1] and R[0, 0] > R[2, 2]:
        s = np.sqrt(1.0 + R[0, 0] - R[1, 1] - R[2, 2]) * 2
        q = [(R[2, 1] - R[1, 2]) / s, 0.25 * s, (R[0, 1] + R[1, 0]) / s, (R[0, 2] + R[2, 0]) / s]
    elif R[1, 1] > R[2, 2]:
        s = np.sqrt(1.0 + R[1, 1] - R[0, 0] - R[2, 2]) * 2
        q = [(R[0, 2] - R[2, 0]) / s, (R[0, 1] + R[1, 0]) / s, 0.25 * s, (R[1, 2] + R[2, 1]) / s]
    else:
        s = np.sqrt(1.0 + R[2, 2] - R[0, 0] - R[1, 1]) * 2
        q = [(R[1, 0] - R[0, 1]) / s, (R[0, 2] + R[2, 0]) / s, (R[1, 2] + R[2, 1]) / s, 0.25 * s]
    q = quat_normalize(q)
    return q if q[0] >= 0 else -q


def euler2quat(e, seq="xyz"):
    """MuJoCo eulerseq semantics: lower case = intrinsic (rotating frame)."""
    q = np.array([1.0, 0, 0, 0])
    for ang, ch in zip(e, seq):
        ax = "xyz".index(ch.lower())
        r = np.zeros(4)
        r[0] = np.cos(ang / 2)
        r[1 + ax] = np.sin(ang / 2)
        q = quat_mul(q, r) if ch.islower() else quat_mul(r, q)
    return quat_normalize(q)


def fl(s, n=None, default=None):
    if s is None:
        return None if default is None else np.array(default, dtype=np.float64)
    v = np.array([float(x) for x in s.split()], dtype=np.float64)
    if n is not None and default is not None and v.size < n:
        full = np.array(default, dtype=np.float64)
        full[: v.size] = v
        v = full
    return v


# ----------------------------------------------------------------------------- meshes
def read_stl(path):
    raw = open(path, "rb").read()
    n = int(np.frombuffer(raw[80:84], dtype=np.uint32)[0])
    rec = np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")])
    data = np.frombuffer(raw[84: 84 + n * rec.itemsize], dtype=rec)
    return data["v"].copy()  # (n,3,3) float32


def process_mesh(path):
    """Legacy (exactmeshinertia=false) mesh frame: CoM + principal axes, float32 verts, hull."""
    from scipy.spatial import ConvexHull

    tris = read_stl(path)
    verts32, inv = np.unique(tris.reshape(-1, 3), axis=0, return_inverse=True)
    faces = inv.reshape(-1, 3)
    v = verts32.astype(np.float64)
    a, b, c = v[faces[:, 0]], v[faces[:, 1]], v[faces[:, 2]]
    nrm = np.cross(b - a, c - a)
    area2 = np.linalg.norm(nrm, axis=1)
    ok = area2 > 1e-20
    a, b, c, nrm, area2 = a[ok], b[ok], c[ok], nrm[ok], area2[ok]
    area = 0.5 * area2
    n = nrm / area2[:, None]
    cen = (a + b + c) / 3.0
    facecen = (area[:, None] * cen).sum(0) / area.sum()
    # pyramids with apex at the area-weighted face centroid, |volume| (legacy: exact for convex)
    vol = np.abs(area * np.einsum("ij,ij->i", cen - facecen, n) / 3.0)
    com = (vol[:, None] * (0.75 * cen + 0.25 * facecen)).sum(0) / vol.sum()
    # second moments about the CoM from the same pyramids
    p0 = np.broadcast_to(facecen - com, a.shape)
    P = [p0, a - com, b - com, c - com]
    S = sum(P)
    C = np.zeros((3, 3))
    for p in P:
        C += np.einsum("i,ij,ik->jk", vol, p, p)
    C += np.einsum("i,ij,ik->jk", vol, S, S)
    C /= 20.0
    inertia = np.trace(C) * np.eye(3) - C
    w, V = np.linalg.eigh(inertia)
    order = np.argsort(-w)
    V = V[:, order]
    for k in range(2):  # canonical sign: largest-magnitude component positive
        j = np.argmax(np.abs(V[:, k]))
        if V[j, k] < 0:
            V[:, k] = -V[:, k]
    V[:, 2] = np.cross(V[:, 0], V[:, 1])
    local32 = ((v - com) @ V).astype(np.float32)
    hull = ConvexHull(local32.astype(np.float64), qhull_options="Qt")
    hv = np.sort(hull.vertices)
    hull_verts = local32[hv]
    rbound = float(np.max(np.linalg.norm(local32.astype(np.float64), axis=1)))
    return {
        "pos": com, "quat": mat2quat(V), "verts": hull_verts, "rbound": rbound,
        "ntri": int(tris.shape[0]), "nvert": int(verts32.shape[0]), "nhull": int(hv.size),
        "volume": float(vol.sum()),
    }


# ----------------------------------------------------------------------------- MJCF subset
class Body:
    def __init__(self):
        self.name = ""
        self.parent = 0
        self.pos = np.zeros(3)
        self.quat = np.array([1.0, 0, 0, 0])
        self.ipos = np.zeros(3)
        self.iquat = np.array([1.0, 0, 0, 0])
        self.mass = 0.0
        self.inertia = np.zeros(3)
        self.explicit_inertial = False
        self.joints = []
        self.geoms = []
        self.sites = []


def load_mjcf(world_xml):
    root = ET.parse(world_xml).getroot()
    base = os.path.dirname(world_xml)
    # resolve top-level includes in place (MuJoCo pastes the included root's children)
    children = []
    for el in list(root):
        if el.tag == "include":
            inc_path = os.path.join(base, el.get("file"))
            inc = ET.parse(inc_path).getroot()
            children.extend([(c, os.path.dirname(inc_path)) for c in inc])
        else:
            children.append((el, base))

    angle_radian = any(c.tag == "compiler" and c.get("angle") == "radian" for c, _ in children)
    assert angle_radian, "this compiler only handles angle=radian"

    dflt = {"geom": {}, "joint": {}, "motor": {}}
    for c, _ in children:
        if c.tag == "default":
            for e in c:
                if e.tag in dflt:
                    dflt[e.tag].update(e.attrib)

    meshes, hfields = {}, {}
    for c, d in children:
        if c.tag == "asset":
            for e in c:
                if e.tag == "mesh":
                    meshes[e.get("name")] = os.path.join(d, e.get("file"))
                elif e.tag == "hfield":
                    hfields[e.get("name")] = e.attrib

    bodies = [Body()]
    bodies[0].name = "world"

    def attr(el, key, tag):
        return el.get(key, dflt.get(tag, {}).get(key))

    def parse_body_children(el, bid):
        for e in el:
            if e.tag == "inertial":
                b = bodies[bid]
                b.explicit_inertial = True
                b.ipos = fl(e.get("pos"), default=[0, 0, 0])
                if e.get("euler") is not None:
                    b.iquat = euler2quat(fl(e.get("euler")))
                else:
                    b.iquat = quat_normalize(fl(e.get("quat"), default=[1, 0, 0, 0]))
                b.mass = float(e.get("mass"))
                b.inertia = fl(e.get("diaginertia"))
            elif e.tag == "joint":
                jt = e.get("type", dflt["joint"].get("type", "hinge"))
                j = {
                    "name": e.get("name"), "type": jt,
                    "pos": fl(attr(e, "pos", "joint"), default=[0, 0, 0]),
                    "axis": fl(attr(e, "axis", "joint"), default=[0, 0, 1]),
                    "range": fl(attr(e, "range", "joint"), default=[0, 0]),
                    "damping": float(attr(e, "damping", "joint") or 0),
                    "stiffness": float(attr(e, "stiffness", "joint") or 0),
                    "armature": float(attr(e, "armature", "joint") or 0),
                    "margin": float(attr(e, "margin", "joint") or 0),
                    "limited": (attr(e, "limited", "joint") == "true") and jt != "free",
                    "solref": fl(attr(e, "solreflimit", "joint"), default=[0.02, 1]),
                    "solimp": fl(attr(e, "solimplimit", "joint"), 5, [0.9, 0.95, 0.001, 0.5, 2]),
                }
                j["axis"] = j["axis"] / np.linalg.norm(j["axis"])
                bodies[bid].joints.append(j)
            elif e.tag == "geom":
                g = {
                    "name": e.get("name"),
                    "type": attr(e, "type", "geom") or "sphere",
                    "pos": fl(e.get("pos"), default=[0, 0, 0]),
                    "quat": quat_normalize(fl(e.get("quat"), default=[1, 0, 0, 0])),
                    "mesh": e.get("mesh"), "hfield": e.get("hfield"),
                    "contype": int(attr(e, "contype", "geom") or 1),
                    "conaffinity": int(attr(e, "conaffinity", "geom") or 1),
                    "condim": int(attr(e, "condim", "geom") or 3),
                    "friction": fl(attr(e, "friction", "geom"), 3, [1, 0.005, 0.0001]),
                    "solref": fl(attr(e, "solref", "geom"), default=[0.02, 1]),
                    "solimp": fl(attr(e, "solimp", "geom"), 5, [0.9, 0.95, 0.001, 0.5, 2]),
                    "solmix": float(attr(e, "solmix", "geom") or 1),
                    "margin": float(attr(e, "margin", "geom") or 0),
                    "gap": float(attr(e, "gap", "geom") or 0),
                }
                bodies[bid].geoms.append(g)
            elif e.tag == "site":
                bodies[bid].sites.append({"name": e.get("name"), "pos": fl(e.get("pos"), default=[0, 0, 0])})
            elif e.tag == "body":
                nb = Body()
                nb.name = e.get("name")
                nb.parent = bid
                nb.pos = fl(e.get("pos"), default=[0, 0, 0])
                nb.quat = quat_normalize(fl(e.get("quat"), default=[1, 0, 0, 0]))
                bodies.append(nb)
                parse_body_children(e, len(bodies) - 1)

    for c, _ in children:
        if c.tag == "worldbody":
            parse_body_children(c, 0)

    excludes = []
    for c, _ in children:
        if c.tag == "contact":
            for e in c:
                if e.tag == "exclude":
                    excludes.append((e.get("body1"), e.get("body2")))

    actuators = []
    for c, _ in children:
        if c.tag == "actuator":
            for e in c:
                assert e.tag == "motor"
                actuators.append({
                    "name": e.get("name"), "joint": e.get("joint"),
                    "gear": fl(attr(e, "gear", "motor"), default=[1])[0],
                    "ctrlrange": fl(attr(e, "ctrlrange", "motor"), default=[0, 0]),
                    "forcerange": fl(attr(e, "forcerange", "motor"), default=[0, 0]),
                    "ctrllimited": attr(e, "ctrllimited", "motor") == "true",
                    "forcelimited": attr(e, "forcelimited", "motor") == "true",
                })
    return bodies, meshes, hfields, excludes, actuators


# ----------------------------------------------------------------------------- model build
def build(ref_root):
    D = DEFINES
    assets = os.path.join(ref_root, "simulation", "assets")
    bodies, meshes, hfields, excludes, actuators = load_mjcf(os.path.join(assets, "world.xml"))
    nbody = len(bodies)
    assert nbody == D["HMJ_NBODY"], nbody
    m = np.zeros((), dtype=MODEL_DTYPE)
    m["magic"], m["version"], m["sizeof_model"] = D["HMJ_MAGIC"], D["HMJ_VERSION"], MODEL_DTYPE.itemsize

    # mjOption defaults (no <option> element in either XML)
    m["timestep"] = 0.002
    m["gravity"] = [0, 0, -9.81]
    m["tolerance"], m["ls_tolerance"], m["mpr_tolerance"] = 1e-8, 0.01, 1e-6
    m["iterations"], m["ls_iterations"], m["mpr_iterations"] = 100, 50, 50

    name2body = {b.name: i for i, b in enumerate(bodies)}
    # ---- bodies / joints / dofs
    jnt, dofs = [], []
    qadr = 0
    for i, b in enumerate(bodies):
        m["body_parentid"][i] = b.parent
        m["body_pos"][i], m["body_quat"][i] = b.pos, b.quat
        m["body_ipos"][i], m["body_iquat"][i] = b.ipos, b.iquat
        m["body_mass"][i], m["body_inertia"][i] = b.mass, b.inertia
        m["body_sameframe"][i] = int(np.all(b.ipos == 0) and np.all(b.iquat == [1, 0, 0, 0]))
        m["body_jntadr"][i] = len(jnt) if b.joints else -1
        m["body_jntnum"][i] = len(b.joints)
        m["body_dofadr"][i] = len(dofs) if b.joints else -1
        nd = 0
        for j in b.joints:
            jid = len(jnt)
            j["body"], j["qposadr"], j["dofadr"] = i, qadr, len(dofs)
            n_q, n_v = (7, 6) if j["type"] == "free" else (1, 1)
            for k in range(n_v):
                dofs.append({"body": i, "jnt": jid})
            qadr += n_q
            nd += n_v
            jnt.append(j)
        m["body_dofnum"][i] = nd
        if i == 0:
            m["body_rootid"][i] = m["body_weldid"][i] = 0
            m["body_depth"][i] = 0
        else:
            m["body_rootid"][i] = i if b.parent == 0 else m["body_rootid"][b.parent]
            m["body_weldid"][i] = i if b.joints else m["body_weldid"][b.parent]
            m["body_depth"][i] = m["body_depth"][b.parent] + 1
    assert len(jnt) == D["HMJ_NJNT"] and len(dofs) == D["HMJ_NV"] and qadr == D["HMJ_NQ"]

    jtype = {"free": D["HMJ_JNT_FREE"], "hinge": D["HMJ_JNT_HINGE"]}
    for jid, j in enumerate(jnt):
        m["jnt_type"][jid] = jtype[j["type"]]
        m["jnt_qposadr"][jid], m["jnt_dofadr"][jid], m["jnt_bodyid"][jid] = j["qposadr"], j["dofadr"], j["body"]
        m["jnt_limited"][jid] = int(j["limited"])
        m["jnt_pos"][jid], m["jnt_axis"][jid] = j["pos"], j["axis"]
        m["jnt_stiffness"][jid] = j["stiffness"]
        m["jnt_range"][jid] = j["range"]
        m["jnt_margin"][jid] = j["margin"]
        m["jnt_solref"][jid], m["jnt_solimp"][jid] = j["solref"], j["solimp"]
        qa = j["qposadr"]
        if j["type"] == "free":
            b = bodies[j["body"]]
            m["qpos0"][qa: qa + 3], m["qpos0"][qa + 3: qa + 7] = b.pos, b.quat
        else:
            m["qpos0"][qa] = 0.0
    m["qpos_spring"] = m["qpos0"]  # springref = 0 for hinges; free joints spring to qpos0

    # dof tables: parent dof = previous dof in the same body, else last dof of the nearest jointed ancestor
    for d, info in enumerate(dofs):
        b = info["body"]
        m["dof_bodyid"][d], m["dof_jntid"][d] = b, info["jnt"]
        if d > m["body_dofadr"][b]:
            m["dof_parentid"][d] = d - 1
        else:
            p = bodies[b].parent
            while p and m["body_dofnum"][p] == 0:
                p = bodies[p].parent
            m["dof_parentid"][d] = (m["body_dofadr"][p] + m["body_dofnum"][p] - 1) if p else -1
        j = jnt[info["jnt"]]
        m["dof_armature"][d], m["dof_damping"][d] = j["armature"], j["damping"]
    adr = 0
    for d in range(len(dofs)):
        m["dof_Madr"][d] = adr
        k = d
        while k >= 0:
            adr += 1
            k = m["dof_parentid"][k]
    m["dof_Madr"][len(dofs)] = adr
    assert adr == D["HMJ_NM"], adr

    # ---- actuators
    name2jnt = {j["name"]: i for i, j in enumerate(jnt)}
    assert len(actuators) == D["HMJ_NU"]
    for a, act in enumerate(actuators):
        jid = name2jnt[act["joint"]]
        assert act["ctrllimited"] and act["forcelimited"]
        m["act_dofadr"][a], m["act_qposadr"][a] = jnt[jid]["dofadr"], jnt[jid]["qposadr"]
        m["act_gear"][a], m["act_gainprm0"][a] = act["gear"], 1.0
        m["act_ctrlrange"][a], m["act_forcerange"][a] = act["ctrlrange"], act["forcerange"]

    # ---- meshes (only those referenced by colliding geoms), geoms
    mesh_info, mesh_ids = {}, {}
    col_geoms = []  # (mujoco geom id, body id, geom dict)
    gid = 0
    for bi, b in enumerate(bodies):
        for g in b.geoms:
            if g["contype"] or g["conaffinity"]:
                col_geoms.append((gid, bi, g))
            gid += 1
    assert len(col_geoms) == D["HMJ_NGEOM"], len(col_geoms)
    assert col_geoms[0][2]["type"] == "hfield"
    vadr = 0
    report = {}
    for k, (mjid, bi, g) in enumerate(col_geoms):
        m["geom_bodyid"][k], m["geom_mjid"][k] = bi, mjid
        m["geom_friction"][k], m["geom_solref"][k] = g["friction"], g["solref"]
        m["geom_solimp"][k], m["geom_solmix"][k] = g["solimp"], g["solmix"]
        assert g["condim"] == 6 and g["margin"] == 0 and g["gap"] == 0
        if g["type"] == "hfield":
            hf = hfields[g["hfield"]]
            assert int(hf["nrow"]) == D["HMJ_HF_NROW"] and int(hf["ncol"]) == D["HMJ_HF_NCOL"]
            m["geom_type"][k], m["geom_meshid"][k] = D["HMJ_GEOM_HFIELD"], -1
            m["hfield_size"] = fl(hf["size"])
            m["hfield_nrow"], m["hfield_ncol"] = int(hf["nrow"]), int(hf["ncol"])
            m["hfield_data"][:] = 0.0  # no file: flat until reset writes per-env data
            m["geom_pos"][k], m["geom_quat"][k] = g["pos"], g["quat"]
            s = m["hfield_size"]
            m["geom_rbound"][k] = float(np.sqrt(s[0] ** 2 + s[1] ** 2 + max(s[2], s[3]) ** 2))
        else:
            assert g["type"] == "mesh"
            name = g["mesh"]
            if name not in mesh_ids:
                info = process_mesh(meshes[name])
                mid = len(mesh_ids)
                mesh_ids[name], mesh_info[name] = mid, info
                nv = info["verts"].shape[0]
                m["mesh_vertadr"][mid], m["mesh_vertnum"][mid] = vadr, nv
                m["mesh_vert"][vadr: vadr + nv] = info["verts"]
                vadr += nv
                report[name] = {k2: info[k2] for k2 in ("ntri", "nvert", "nhull", "volume", "rbound")}
            info = mesh_info[name]
            m["geom_type"][k], m["geom_meshid"][k] = D["HMJ_GEOM_MESH"], mesh_ids[name]
            m["geom_pos"][k] = g["pos"] + quat2mat(g["quat"]) @ info["pos"]
            m["geom_quat"][k] = quat_normalize(quat_mul(g["quat"], info["quat"]))
            m["geom_rbound"][k] = info["rbound"]
    assert len(mesh_ids) == D["HMJ_NMESH"], len(mesh_ids)
    assert vadr <= D["HMJ_MAXVERT"], vadr

    # ---- sites
    sid = 0
    for bi, b in enumerate(bodies):
        for s in b.sites:
            m["site_bodyid"][sid], m["site_pos"][sid] = bi, s["pos"]
            sid += 1
    assert sid == D["HMJ_NSITE"]

    # ---- candidate collision pairs in contact order (body-pair signature (b1<<16)+b2)
    excl = set()
    for n1, n2 in excludes:
        b1, b2 = name2body[n1], name2body[n2]
        excl.add((min(b1, b2), max(b1, b2)))
    weld = m["body_weldid"]
    par = m["body_parentid"]

    def filtered(b1, b2):
        w1, w2 = weld[b1], weld[b2]
        wp1, wp2 = weld[par[w1]], weld[par[w2]]
        if w1 == w2:
            return True
        if w1 != 0 and w2 != 0 and (w1 == wp2 or w2 == wp1):
            return True
        return (min(b1, b2), max(b1, b2)) in excl

    pairs = []
    for k1 in range(len(col_geoms)):
        for k2 in range(k1 + 1, len(col_geoms)):
            b1, b2 = col_geoms[k1][1], col_geoms[k2][1]
            if b1 == b2 or filtered(b1, b2):
                continue
            pairs.append(((min(b1, b2) << 16) + max(b1, b2), k1, k2))
    pairs.sort()
    assert len(pairs) == D["HMJ_NPAIR"], len(pairs)
    m["npair"] = len(pairs)
    for p, (_, k1, k2) in enumerate(pairs):
        m["pair_geom1"][p], m["pair_geom2"][p] = k1, k2

    # ---- constants at qpos0 (mj_setConst): subtree mass, invweight0, meaninertia
    subtree = np.array([b.mass for b in bodies])
    for i in range(nbody - 1, 0, -1):
        subtree[bodies[i].parent] += subtree[i]
    m["body_subtreemass"] = subtree

    xpos = np.zeros((nbody, 3))
    xquat = np.zeros((nbody, 4))
    xquat[0] = [1, 0, 0, 0]
    for i in range(1, nbody):
        b = bodies[i]
        R = quat2mat(xquat[b.parent])
        xpos[i] = xpos[b.parent] + R @ b.pos
        xquat[i] = quat_normalize(quat_mul(xquat[b.parent], b.quat))  # qpos == qpos0: hinge angles 0
    nv = len(dofs)
    Jp = np.zeros((nbody, 3, nv))
    Jr = np.zeros((nbody, 3, nv))
    xipos = np.zeros((nbody, 3))
    M = np.diag(np.array(m["dof_armature"], dtype=np.float64))
    for i in range(1, nbody):
        Ri = quat2mat(xquat[i])
        xipos[i] = xpos[i] + Ri @ bodies[i].ipos
        # walk up the dof chain
        b = i
        while b:
            for j in bodies[b].joints:
                Rb = quat2mat(xquat[b])
                if j["type"] == "free":
                    for k in range(3):
                        Jp[i, k, j["dofadr"] + k] = 1.0
                        ax = Rb[:, k]
                        Jr[i, :, j["dofadr"] + 3 + k] = ax
                        Jp[i, :, j["dofadr"] + 3 + k] = np.cross(ax, xipos[i] - xpos[b])
                else:
                    ax = Rb @ j["axis"]
                    anchor = xpos[b] + Rb @ j["pos"]
                    Jr[i, :, j["dofadr"]] = ax
                    Jp[i, :, j["dofadr"]] = np.cross(ax, xipos[i] - anchor)
            b = bodies[b].parent
        Rin = Ri @ quat2mat(bodies[i].iquat)
        Iw = Rin @ np.diag(bodies[i].inertia) @ Rin.T
        M += bodies[i].mass * Jp[i].T @ Jp[i] + Jr[i].T @ Iw @ Jr[i]
    A = np.linalg.inv(M)
    m["meaninertia"] = float(np.mean(np.diag(M)))
    for jid, j in enumerate(jnt):
        d0 = j["dofadr"]
        if j["type"] == "free":
            m["dof_invweight0"][d0: d0 + 3] = np.mean(np.diag(A)[d0: d0 + 3])
            m["dof_invweight0"][d0 + 3: d0 + 6] = np.mean(np.diag(A)[d0 + 3: d0 + 6])
        else:
            m["dof_invweight0"][d0] = A[d0, d0]
    for i in range(1, nbody):
        m["body_invweight0"][i, 0] = np.trace(Jp[i] @ A @ Jp[i].T) / 3.0
        m["body_invweight0"][i, 1] = np.trace(Jr[i] @ A @ Jr[i].T) / 3.0

    # ---- planner task (task.xml): cost terms from leading <user> sensors, customs
    task_xml = os.path.join(ref_root, "mujoco_mpc", "mjpc", "tasks", "humanoid_cap", "stand", "task.xml")
    troot = ET.parse(task_xml).getroot()
    customs = {e.get("name"): fl(e.get("data")) for e in troot.iter("numeric")}
    terms = []
    for e in troot.find("sensor"):
        if e.tag != "user":
            break
        terms.append((e.get("name"), int(e.get("dim")), fl(e.get("user"))))
    assert len(terms) == D["HMJ_NTERM"]
    nparam = {-1: 0, 0: 0, 1: 2, 2: 1, 3: 1, 5: 1, 6: 1, 7: 2, 8: 1}
    m["task_nterm"] = len(terms)
    m["task_nresidual"] = sum(t[1] for t in terms)
    assert m["task_nresidual"] == D["HMJ_NRESIDUAL"]
    for k, (_, dim, user) in enumerate(terms):
        m["task_dim"][k], m["task_norm"][k], m["task_weight"][k] = dim, int(user[0]), user[1]
        npm = nparam[int(user[0])]
        m["task_normprm"][k, :npm] = user[4: 4 + npm]
    m["task_risk"] = float(customs.get("task_risk", [0.0])[0])
    m["task_height_goal"] = customs["residual_Height Goal"][0]
    m["agent_timestep"] = customs["agent_timestep"][0]
    m["agent_horizon"] = customs["agent_horizon"][0]
    m["sampling_exploration"] = customs["sampling_exploration"][0]
    m["sampling_spline_points"] = int(customs["sampling_spline_points"][0])
    m["sampling_trajectories"] = int(customs.get("sampling_trajectories", [10])[0])

    report_all = {
        "bodies": [b.name for b in bodies],
        "joints": [j["name"] for j in jnt],
        "actuators": [a["name"] for a in actuators],
        "col_geoms": [{"mjid": g[0], "body": bodies[g[1]].name, "name": g[2]["name"], "mesh": g[2]["mesh"]} for g in col_geoms],
        "meshes": report,
        "pairs": [[int(k1), int(k2)] for _, k1, k2 in pairs],
        "nvert_total": int(vadr),
    }
    return m, report_all


def to_jsonable(m):
    out = {}
    for name in m.dtype.names:
        v = m[name]
        if name == "mesh_vert":
            continue
        out[name] = v.tolist() if isinstance(v, np.ndarray) else v.item()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default=DEFAULT_REF)
    ap.add_argument("--out", default=os.path.join(REPO, "humanoid-mujoco_b200", "assets"))
    args = ap.parse_args()
    m, report = build(args.ref)
    os.makedirs(args.out, exist_ok=True)
    m.tofile(os.path.join(args.out, "humanoid_model.bin"))
    with open(os.path.join(args.out, "humanoid_model.json"), "w") as f:
        json.dump({"report": report, "model": to_jsonable(m)}, f, indent=1)
    print(json.dumps(report, indent=1))
    print("sizeof(hmj_model_t) =", MODEL_DTYPE.itemsize, " meaninertia =", float(m["meaninertia"]))


if __name__ == "__main__":
    main()
