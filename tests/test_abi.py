"""The C-ABI boundary: the product library loads without a GPU, exports every symbol the headers declare, and the
struct layouts match the reference's."""
import ctypes
import os
import re
import subprocess

import pytest

from quest_b200 import capi

ROOT = capi.REPO_ROOT


def _declared_symbols():
    names = set()
    for hdr in ("QuEST.h", "QuEST_debug.h", "quest_b200.h"):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        for m in re.finditer(r"^[A-Za-z_][\w\s\*]*?\b(\w+)\s*\([^;{]*\)\s*;", text, flags=re.M):
            names.add(m.group(1))
    return names


@pytest.fixture(scope="module")
def product_path():
    if not os.path.isfile(capi.PRODUCT_LIB):
        from quest_b200 import build
        build.build_product(2)
    return capi.PRODUCT_LIB


def test_headers_declare_expected_api():
    names = _declared_symbols()
    assert len(names) >= 85  # 73 QuEST.h prototypes + debug API + extensions
    for n in ("createQureg", "hadamard", "multiControlledUnitary", "calcPurity", "mixDamping", "qb200_getStream"):
        assert n in names


def test_product_exports_every_declared_symbol(product_path):
    out = subprocess.run(["nm", "-D", "--defined-only", product_path], capture_output=True, text=True, check=True).stdout
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    missing = sorted(_declared_symbols() - exported)
    assert not missing, missing
    # mt19937ar symbols the reference suite calls through the library
    for n in ("genrand_real1", "init_by_array", "init_genrand"):
        assert n in exported


def test_product_loads_without_gpu_and_reports_precision(product_path):
    lib = capi.load(product_path)  # dlopen only; no CUDA call is made
    assert lib.prec == 2 and not lib.missing


def test_product_does_not_link_oracle_or_openmp(product_path):
    out = subprocess.run(["ldd", product_path], capture_output=True, text=True, check=True).stdout
    assert "oracle" not in out and "gomp" not in out


def test_struct_layout_matches_reference_abi(oracle):
    t = oracle
    assert ctypes.sizeof(t.Qureg) == 112
    offs = {f[0]: getattr(t.Qureg, f[0]).offset for f in t.Qureg._fields_}
    assert offs == {"isDensityMatrix": 0, "numQubitsRepresented": 4, "numQubitsInStateVec": 8, "numAmpsPerChunk": 16,
                    "numAmpsTotal": 24, "chunkId": 32, "numChunks": 36, "stateVec": 40, "pairStateVec": 56,
                    "deviceStateVec": 72, "firstLevelReduction": 88, "secondLevelReduction": 96, "qasmLog": 104}
    assert ctypes.sizeof(t.QuESTEnv) == 8 and ctypes.sizeof(t.Complex) == 16
    assert ctypes.sizeof(t.ComplexMatrix2) == 64 and ctypes.sizeof(t.Vector) == 24 and ctypes.sizeof(t.QASMLogger) == 24


def test_c_struct_sizes_compile_time(tmp_path):
    """The same facts checked by the C compiler against include/QuEST.h."""
    src = tmp_path / "abi.c"
    src.write_text('#include <stddef.h>\n#include "QuEST.h"\n'
                   '_Static_assert(sizeof(Qureg)==112,"Qureg");_Static_assert(offsetof(Qureg,qasmLog)==104,"qasm");\n'
                   '_Static_assert(offsetof(Qureg,deviceStateVec)==72,"dev");_Static_assert(sizeof(QuESTEnv)==8,"env");\n'
                   '_Static_assert(sizeof(ComplexMatrix2)==64,"m2");int main(void){return 0;}\n')
    subprocess.run(["gcc", "-std=c11", "-I", os.path.join(ROOT, "include"), "-c", str(src), "-o", str(tmp_path / "abi.o")], check=True)


def test_validation_exits_with_reference_code_and_text(oracle):
    """Invalid input prints the reference's message and exits with its enum code (QuEST_validation.c:82-88)."""
    code = ("import sys; sys.path.insert(0, %r)\nfrom quest_b200 import capi\nlib = capi.load(capi.ORACLE_LIB)\n"
            "env = lib.createQuESTEnv(); q = lib.createQureg(3, env); lib.hadamard(q, 5)\n" % ROOT)
    r = subprocess.run(["python", "-c", code], capture_output=True, text=True)
    assert r.returncode == 2  # E_INVALID_TARGET_QUBIT
    assert "QuEST Error in function hadamard: Invalid target qubit. Note qubits are zero indexed." in r.stdout
    code2 = code.replace("lib.hadamard(q, 5)", "lib.controlledNot(q, 1, 1)")
    r = subprocess.run(["python", "-c", code2], capture_output=True, text=True)
    assert r.returncode == 7 and "Control qubit cannot equal target qubit." in r.stdout
