"""The 4-wide traversal tree (csrc/flat.h nodes4, built by host_flatten.cpp collapse4): below the scene root, every
volume boundary and every object entry it must reach exactly the leaf slots of the 2-wide traversal tree, each with the
same gate box, and carry no interior box the 2-wide tree does not hold — which is why wf_extend's 4-wide walk returns
the same hits (tests/test_gpu_wavefront.py checks that on the device).  Host code only."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(ROOT, "raytracingrust_b200", "csrc")


def test_four_wide_tree_reaches_the_same_gated_leaves(tmp_path):
    exe = str(tmp_path / "check_nodes4")
    subprocess.run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-pthread", "-I", os.path.join(ROOT, "include"), "-I", CSRC,
                    os.path.join(HERE, "csrc", "check_nodes4.cpp"), os.path.join(CSRC, "host_flatten.cpp"), "-o", exe], check=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.startswith("OK"), r.stdout[-2000:]
