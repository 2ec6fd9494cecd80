"""The C ABI: ctypes mirror vs include/rtr.h (sizeof/offsetof from a compiled C probe), the built
library exports every declared symbol, and the product fails loudly without a GPU."""
import ctypes as C
import os
import re
import subprocess

import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rtr.h")

STRUCTS = {
    "rtr_vec3": A.Vec3, "rtr_camera": A.CameraS, "rtr_settings": A.SettingsS, "rtr_object_ref": A.ObjectRef,
    "rtr_sphere": A.SphereS, "rtr_mesh": A.MeshS, "rtr_volume": A.VolumeS, "rtr_material": A.MaterialS,
    "rtr_background": A.BackgroundS, "rtr_scene_desc": A.SceneDesc, "rtr_render_params": A.RenderParams,
    "rtr_stats": A.Stats, "rtr_ray": A.RayS, "rtr_hit": A.HitS, "rtr_scatter_in": A.ScatterIn,
    "rtr_scatter_out": A.ScatterOut,
}


def test_struct_layouts_match_header(tmp_path):
    lines = ['#include <stdio.h>', '#include <stddef.h>', f'#include "{HEADER}"', 'int main(void){']
    for cname, ct in STRUCTS.items():
        lines.append(f'printf("{cname} %zu\\n", sizeof({cname}));')
        for fname, _ in ct._fields_:
            lines.append(f'printf("{cname}.{fname} %zu\\n", offsetof({cname}, {fname}));')
    lines.append('return 0;}')
    src = tmp_path / "probe.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "probe"
    subprocess.check_call(["/usr/bin/gcc", "-std=c11", "-Wall", "-Werror", str(src), "-o", str(exe)])
    got = dict(l.split() for l in subprocess.check_output([str(exe)], text=True).splitlines())
    for cname, ct in STRUCTS.items():
        assert int(got[cname]) == C.sizeof(ct), cname
        for fname, _ in ct._fields_:
            assert int(got[f"{cname}.{fname}"]) == getattr(ct, fname).offset, f"{cname}.{fname}"


def test_header_is_plain_c(tmp_path):
    src = tmp_path / "c89.c"
    src.write_text(f'#include "{HEADER}"\nint main(void){{return (int)RTR_ABI_VERSION - 1;}}\n')
    subprocess.check_call(["/usr/bin/gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", str(src), "-o", str(tmp_path / "a")])


def test_declared_symbols_match_list():
    text = open(HEADER).read()
    declared = set(re.findall(r"\b(rtr_[a-z_0-9]+)\s*\(", text))
    assert declared == set(A.EXPORTED_SYMBOLS)


def test_library_exports_every_symbol():
    L = rt.lib()  # loads librtr_b200.so; no compute
    for sym in A.EXPORTED_SYMBOLS:
        assert hasattr(L, sym), sym
    assert L.rtr_abi_version() == A.RTR_ABI_VERSION


def test_no_cpu_fallback():
    """Without a CUDA device the product must refuse, not fall back (there is no CPU path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(rt.RtrError) as e:
        rt.Context(0)
    assert e.value.status == A.RTR_CUDA_ERROR
    assert "no CPU fallback" in str(e.value)


def test_product_does_not_reference_oracle():
    pkg = os.path.join(ROOT, "raytracingrust_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "liboracle" not in text and "import oracle" not in text and "from oracle" not in text, f
                assert "oracle/" not in text, f
