"""The streaming kernel divides by constants (1000 ms/s, the window, the sampled interval) with
the three-operation Markstein sequence instead of the division routine (csrc/cg_math.cuh
div_const).  That is only legitimate if it returns the correctly rounded quotient: check it on
the host against IEEE division for the constants in use."""
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))


def test_markstein_constant_division_is_exact():
    with tempfile.TemporaryDirectory() as d:
        exe = os.path.join(d, "div_const_check")
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-mfma", "-o", exe,
                               os.path.join(HERE, "div_const_check.c"), "-lm"])
        out = subprocess.check_output([exe], text=True).strip()
    assert out == "0", f"{out} quotients differ from IEEE division"
