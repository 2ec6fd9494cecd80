"""sphere_t's exact shortcuts (csrc/device_core.cuh) keep Sphere::hit (lib/objects.rs:23-44) bit for bit: CPU sweep
of the reference sequence against the shortcut form over random, on-surface (c == 0) and adversarial
(inf / NaN / denormal / overflow) inputs.  The GPU parity tests cover the device code itself."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def test_shortcut_form_is_bit_identical_to_the_reference_sequence(tmp_path):
    exe = str(tmp_path / "sphere_shortcuts")
    subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-o", exe, os.path.join(HERE, "csrc", "sphere_shortcuts.c"), "-lm"],
                   check=True)
    r = subprocess.run([exe, "30000000"], capture_output=True, text=True)
    assert r.returncode == 0 and "mismatches 0" in r.stdout, r.stdout[-2000:]
