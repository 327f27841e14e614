"""The reference's own acceptance program, unchanged: /root/reference/tutorial_example.c (its CMake default USER_SOURCE,
CMakeLists.txt:11) - 30 qubits, 490 gate calls of 17 kinds, then calcProbOfOutcome on every qubit and getAmp on ten
amplitudes, written to probs.dat / stateVector.dat.

oracle/Makefile (`make tutorial`, run by __graft_entry__.build()) compiles that file where it lies, once against the
reference CPU/OpenMP library (oracle/_ref/tutorial_ref) and once against this repo's headers and libQuEST.so
(oracle/_ref/tutorial_b200); the binaries travel to the GPU box with the rest of oracle/_ref.  Here both run and must
write the same two files, character for character (the program prints six decimals).
"""
import os
import re
import subprocess
import time

import pytest

from quest_b200 import capi

REF_DIR = os.path.dirname(capi.REFERENCE_LIB)


def _run(exe, cwd, env=None):
    t0 = time.perf_counter()
    out = subprocess.run([exe], cwd=cwd, capture_output=True, text=True, timeout=1500, env=env)
    dt = time.perf_counter() - t0
    assert out.returncode == 0, out.stdout + out.stderr
    files = {name: open(os.path.join(cwd, name)).read() for name in ("probs.dat", "stateVector.dat")}
    said = re.search(r"takes time\s+([0-9.]+) seconds", out.stdout)
    return files, float(said.group(1)) if said else dt


@pytest.mark.gpu
def test_reference_tutorial_program_relinked_unchanged(tmp_path):
    ref_exe, gpu_exe = os.path.join(REF_DIR, "tutorial_ref"), os.path.join(REF_DIR, "tutorial_b200")
    if not (os.path.isfile(ref_exe) and os.path.isfile(gpu_exe)):
        pytest.skip("oracle/_ref/tutorial_{ref,b200} not built (needs /root/reference at build time)")
    (tmp_path / "gpu").mkdir()
    (tmp_path / "cpu").mkdir()
    gpu_files, gpu_s = _run(gpu_exe, str(tmp_path / "gpu"))
    cpu_env = dict(os.environ, OMP_NUM_THREADS=str(os.cpu_count() or 1))
    cpu_files, cpu_s = _run(ref_exe, str(tmp_path / "cpu"), env=cpu_env)
    print(f"tutorial_example.c: quest-b200 {gpu_s:.2f} s, reference CPU/OpenMP build {cpu_s:.1f} s "
          f"({os.cpu_count()} host threads); the file's own header comment estimates 3783.9 s")
    assert len(re.findall(r"Probability for q\[", gpu_files["probs.dat"])) == 30
    assert len(re.findall(r"Amplitude of", gpu_files["stateVector.dat"])) == 10
    assert gpu_files["probs.dat"] == cpu_files["probs.dat"]
    assert gpu_files["stateVector.dat"] == cpu_files["stateVector.dat"]
    with open(os.path.join(capi.REPO_ROOT, "gpurun_out", "tutorial_times.txt"), "w") if os.path.isdir(
            os.path.join(capi.REPO_ROOT, "gpurun_out")) else open(os.devnull, "w") as f:
        f.write(f"gpu_s {gpu_s:.3f}\ncpu_s {cpu_s:.3f}\ncpu_threads {os.cpu_count()}\n")
