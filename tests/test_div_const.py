"""The streaming kernel replaces the generic ~45-instruction IEEE division routine by two short
sequences (csrc/cg_math.cuh): div_const (Markstein, for the constants 1000 ms/s, the window and
the sampled interval) and ddiv_pos (float reciprocal seed, two Newton steps, Markstein quotient
with two residual corrections; operands outside a safe range take the true division).  That is
only legitimate if both return the correctly rounded quotient, because decisions must be
bit-identical to the reference's Python floats: tests/div_check.cpp compiles the functions
themselves as plain C++ and compares them with IEEE division on random, integer-valued and
adversarial operands (all-ones divisor significands, quotients next to a rounding boundary,
zeros, denormals, infinities, NaN).  3.2e8 + 2.6e8 operand pairs were checked once offline
(`div_check 400`); this test keeps a 4-million-per-class version."""
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))


def test_division_sequences_are_correctly_rounded():
    with tempfile.TemporaryDirectory() as d:
        exe = os.path.join(d, "div_check")
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-mfma", "-std=c++17", "-Wno-unknown-pragmas",
                               "-o", exe, os.path.join(HERE, "div_check.cpp")])
        out = subprocess.check_output([exe, "4"], text=True).split()
    assert out == ["0", "0"], f"div_const / ddiv_pos quotients differing from IEEE division: {out}"
