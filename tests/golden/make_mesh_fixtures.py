"""Generates tests/golden/meshes/*.npz and tests/golden/scenes/*.json from the reference checkout.

Run once in the build container (where /root/reference exists):  python tests/golden/make_mesh_fixtures.py
The GPU box has no /root/reference; tests and bench.py read only these committed fixtures.
Meshes are stored as the (positions, normals, indices) buffers Mesh::from_file produces
(mesh.rs:63-76), parsed by tests/_pyload.parse_obj; scene JSONs are byte-for-byte the shipped
fixtures (data, not source).
"""
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
from _pyload import parse_obj  # noqa: E402

REF = "/root/reference"
os.makedirs(os.path.join(HERE, "meshes"), exist_ok=True)
os.makedirs(os.path.join(HERE, "scenes"), exist_ok=True)
for name in ["monkey", "fancy_monkey", "house", "plane", "donut", "cylinder", "cube"]:
    p, n, i = parse_obj(os.path.join(REF, "assets", name + ".obj"))
    np.savez_compressed(os.path.join(HERE, "meshes", name + ".npz"), positions=p, normals=n, indices=i)
    print(name, len(p), "vertices", len(i) // 3, "triangles")
for name in ["simple.json", "benchmark.json"]:
    shutil.copyfile(os.path.join(REF, "scenes", name), os.path.join(HERE, "scenes", name))
