"""Independent (pure Python) readers of the reference's on-disk formats, used to feed the oracle
and to cross-check the library's C++ loaders.  Test infrastructure only."""
import json
import os

import numpy as np

import raytracingrust_b200 as rt

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
f32 = np.float32


def _v(d):
    return rt.Vec3(f32(d["x"]), f32(d["y"]), f32(d["z"]))


def _c(d):
    return rt.Color(float(f32(d["r"])), float(f32(d["g"])), float(f32(d["b"])))


def _material(m):
    t = m["type"]
    if t == "Lambertian":
        return rt.Lambertian(_c(m["albedo"]))
    if t == "Metal":
        return rt.Metal(_c(m["albedo"]), float(f32(m["fuzz"])))
    if t == "Emission":
        return rt.Emission(_c(m["color"]), 1.0)
    if t == "Dielectric":
        return rt.Dielectric(float(f32(m["ir"])))
    if t == "Isotropic":
        return rt.Isotropic(_c(m["color"]))
    if t == "MixMaterial":
        return rt.MixMaterial(_material(m["first"]), _material(m["second"]), float(f32(m["factor"])))
    raise ValueError(t)


def _object(o):
    t = o["type"]
    if t == "Sphere":
        return rt.Sphere(_v(o["center"]), float(f32(o["radius"])), rt.MaterialHandle(o["material"]))
    if t == "Volume":
        v = rt.Volume(_object(o["boundary"]), 1.0)
        v.neg_inv_density = float(f32(o["neg_inv_density"]))
        return v
    raise ValueError(t)


def scene_from_json(path):
    """serde/typetag scene JSON (scene.rs:74-76) parsed in Python; clamp_indirect defaults to 10."""
    d = json.load(open(path))
    c, s = d["camera"], d["settings"]
    cam = rt.Camera(_v(c["lookfrom"]), _v(c["lookat"]), _v(c["vertical"]), float(f32(c["vertical_fov"])),
                    float(f32(c["aspect_ratio"])))
    st = rt.RenderSettings(s["samples_per_pixel"], s["max_ray_depth"], float(f32(s.get("clamp_indirect", 10.0))),
                           s["enable_multithreading"], s["enable_bvh_tree"], rt.RenderMode._names[s["mode"]])
    b = d["background"]
    if b["type"] == "UniformBackground":
        bg = rt.UniformBackground(_c(b["color"]))
    elif b["type"] == "GradientBackground":
        bg = rt.GradientBackground(_c(b["top"]), _c(b["bottom"]))
    else:
        img = np.array([[p["r"], p["g"], p["b"]] for p in b["image"]], f32)
        bg = rt.SkyMap(img, b["width"], b["height"])
    sc = rt.Scene(st, cam, bg)
    for m in d["materials"]:
        sc.add_material(_material(m))
    for o in d["objects"]:
        sc.add_object(_object(o))
    return sc


def parse_obj(path):
    """obj-rs load_obj::<Vertex,u32>: triangles with normals, (pos,normal) pairs de-duplicated."""
    pos, nrm, cache, vp, vn, idx = [], [], {}, [], [], []
    for line in open(path):
        t = line.split()
        if not t:
            continue
        if t[0] == "v":
            pos.append([f32(x) for x in t[1:4]])
        elif t[0] == "vn":
            nrm.append([f32(x) for x in t[1:4]])
        elif t[0] == "f":
            if len(t) != 4:
                raise ValueError("Model should be triangulated first to be loaded properly")
            for corner in t[1:]:
                parts = corner.split("/")
                pi, ni = int(parts[0]), int(parts[2])
                pi = pi - 1 if pi > 0 else len(pos) + pi
                ni = ni - 1 if ni > 0 else len(nrm) + ni
                if (pi, ni) not in cache:
                    cache[(pi, ni)] = len(vp)
                    vp.append(pos[pi])
                    vn.append(nrm[ni])
                idx.append(cache[(pi, ni)])
    return np.array(vp, f32).reshape(-1, 3), np.array(vn, f32).reshape(-1, 3), np.array(idx, np.uint32)


def mesh_fixture(name):
    """(positions, normals, indices) of a reference asset, from tests/golden/meshes/<name>.npz."""
    z = np.load(os.path.join(GOLDEN, "meshes", name + ".npz"))
    return z["positions"], z["normals"], z["indices"]
