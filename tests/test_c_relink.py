"""The drop-in boundary seen from C: one unmodified C program (examples/bell_and_sweep.c, written against QuEST.h
only) is compiled once against the reference's own CPU library and once against quest-b200's libQuEST.so; both
binaries must print the same amplitudes, probabilities, measurement outcomes, purity and trace."""
import os
import re
import subprocess

import pytest

from quest_b200 import capi

ROOT = capi.REPO_ROOT
SRC = os.path.join(ROOT, "examples", "bell_and_sweep.c")


def _build(tmp_path, libdir, libname, out):
    exe = str(tmp_path / out)
    # -l:<file> links the given file name, so the same command serves libQuEST.so and libQuEST_oracle.so
    subprocess.run(["gcc", "-O2", "-I", os.path.join(ROOT, "include"), SRC, "-o", exe, "-L", libdir, f"-l:{libname}", "-lm",
                    f"-Wl,-rpath,{libdir}"], check=True)
    return exe


def _numbers(text):
    """Every printed number except the timing line."""
    lines = [ln for ln in text.splitlines() if not ln.startswith("sweep:")]
    sweep = [ln for ln in text.splitlines() if ln.startswith("sweep:")]
    vals = [float(x) for ln in lines for x in re.findall(r"-?\d+\.\d+|-?\b\d+\b", ln)]
    prob = float(re.search(r"total probability = ([0-9.]+)", sweep[0]).group(1))
    return vals, prob


def test_c_program_compiles_against_our_headers_and_cpu_library(tmp_path):
    """CPU side of the relink test: the example builds against include/QuEST.h and runs on a CPU library."""
    if os.path.isfile(capi.REFERENCE_LIB):
        libdir, name = os.path.dirname(capi.REFERENCE_LIB), "libQuEST.so"
    else:
        libdir, name = os.path.dirname(capi.ORACLE_LIB), "libQuEST_oracle.so"
    exe = _build(tmp_path, libdir, name, "demo_cpu")
    out = subprocess.run([exe, "10"], capture_output=True, text=True, check=True).stdout
    assert "GHZ: amp[0] = 0.707106781187" in out and "purity" in out


@pytest.mark.gpu
def test_same_c_program_same_numbers_on_gpu_library(tmp_path):
    if os.path.isfile(capi.REFERENCE_LIB):
        cpu = _build(tmp_path, os.path.dirname(capi.REFERENCE_LIB), "libQuEST.so", "demo_ref")
    else:
        cpu = _build(tmp_path, os.path.dirname(capi.ORACLE_LIB), "libQuEST_oracle.so", "demo_oracle")
    gpu = _build(tmp_path, os.path.dirname(capi.PRODUCT_LIB), "libQuEST.so", "demo_b200")
    for n in ("9", "16"):
        a = subprocess.run([cpu, n], capture_output=True, text=True, check=True).stdout
        b = subprocess.run([gpu, n], capture_output=True, text=True, check=True).stdout
        va, pa = _numbers(a)
        vb, pb = _numbers(b)
        assert va == vb, (a, b)  # printed to 12 decimals: identical text
        assert abs(pa - pb) < 1e-12
