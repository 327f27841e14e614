"""ctypes mirror of the QuEST C API (include/QuEST.h, include/QuEST_debug.h, include/quest_b200.h).

Plays the role the reference's ``utilities/QuESTPy`` package plays for its test-suite
(QuESTPy/QuESTTypes.py:123-137 mirrors the structs by value, QuESTPy/QuESTFunc.py:13-108 declares every entry
point): any shared library that exports the QuEST.h symbols with the reference struct layout can be opened
with :class:`QuESTLibrary` — the product (``quest_b200/lib/libQuEST.so``, CUDA, needs a B200), our CPU oracle
(``oracle/_build/libQuEST_oracle.so``) or the reference's own CPU build (``oracle/_ref/libQuEST.so``) — and
driven with identical calls.  That is what the parity tests do.

Nothing here computes on amplitudes; it only marshals arguments.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Iterable, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)

PRODUCT_LIB = os.path.join(_HERE, "lib", "libQuEST.so")
PRODUCT_LIB_F32 = os.path.join(_HERE, "lib", "libQuEST_f32.so")
ORACLE_LIB = os.path.join(REPO_ROOT, "oracle", "_build", "libQuEST_oracle.so")
ORACLE_LIB_F32 = os.path.join(REPO_ROOT, "oracle", "_build", "libQuEST_oracle_f32.so")
REFERENCE_LIB = os.path.join(REPO_ROOT, "oracle", "_ref", "libQuEST.so")
REFERENCE_LIB_F32 = os.path.join(REPO_ROOT, "oracle", "_ref", "libQuEST_f32.so")


def _make_types(qreal):
    """Struct mirrors for one precision (field order is the ABI, QuEST.h:28-121 in the reference)."""

    class QASMLogger(C.Structure):
        _fields_ = [("buffer", C.c_char_p), ("bufferSize", C.c_int), ("bufferFill", C.c_int), ("isLogging", C.c_int)]

    class ComplexArray(C.Structure):
        _fields_ = [("real", C.POINTER(qreal)), ("imag", C.POINTER(qreal))]

    class Complex(C.Structure):
        _fields_ = [("real", qreal), ("imag", qreal)]

        def __complex__(self):
            return complex(self.real, self.imag)

    class ComplexMatrix2(C.Structure):
        _fields_ = [("r0c0", Complex), ("r0c1", Complex), ("r1c0", Complex), ("r1c1", Complex)]

    class Vector(C.Structure):
        _fields_ = [("x", qreal), ("y", qreal), ("z", qreal)]

    class Qureg(C.Structure):
        _fields_ = [
            ("isDensityMatrix", C.c_int),
            ("numQubitsRepresented", C.c_int),
            ("numQubitsInStateVec", C.c_int),
            ("numAmpsPerChunk", C.c_longlong),
            ("numAmpsTotal", C.c_longlong),
            ("chunkId", C.c_int),
            ("numChunks", C.c_int),
            ("stateVec", ComplexArray),
            ("pairStateVec", ComplexArray),
            ("deviceStateVec", ComplexArray),
            ("firstLevelReduction", C.POINTER(qreal)),
            ("secondLevelReduction", C.POINTER(qreal)),
            ("qasmLog", C.POINTER(QASMLogger)),
        ]

    class QuESTEnv(C.Structure):
        _fields_ = [("rank", C.c_int), ("numRanks", C.c_int)]

    return dict(QASMLogger=QASMLogger, ComplexArray=ComplexArray, Complex=Complex, ComplexMatrix2=ComplexMatrix2,
                Vector=Vector, Qureg=Qureg, QuESTEnv=QuESTEnv)


# name -> (restype, argtypes); tokens are resolved per precision in QuESTLibrary._declare
_Q, _E, _I, _LL, _R, _PR, _PI, _CX, _M2, _V = "Qureg", "QuESTEnv", "int", "ll", "qreal", "pqreal", "pint", "Complex", "ComplexMatrix2", "Vector"
API = {
    # environment / registers
    "createQuESTEnv": (_E, []), "destroyQuESTEnv": (None, [_E]), "syncQuESTEnv": (None, [_E]),
    "syncQuESTSuccess": (_I, [_I]), "reportQuESTEnv": (None, [_E]),
    "createQureg": (_Q, [_I, _E]), "createDensityQureg": (_Q, [_I, _E]), "destroyQureg": (None, [_Q, _E]),
    "cloneQureg": (None, [_Q, _Q]), "reportState": (None, [_Q]), "reportStateToScreen": (None, [_Q, _E, _I]),
    "reportQuregParams": (None, [_Q]), "getNumQubits": (_I, [_Q]), "getNumAmps": (_I, [_Q]),
    "seedQuEST": (None, ["pulong", _I]), "seedQuESTDefault": (None, []), "getQuEST_PREC": (_I, []),
    # QASM
    "startRecordingQASM": (None, [_Q]), "stopRecordingQASM": (None, [_Q]), "clearRecordedQASM": (None, [_Q]),
    "printRecordedQASM": (None, [_Q]), "writeRecordedQASMToFile": (None, [_Q, "cstr"]),
    # initialisation
    "initZeroState": (None, [_Q]), "initPlusState": (None, [_Q]), "initClassicalState": (None, [_Q, _LL]),
    "initPureState": (None, [_Q, _Q]), "initStateFromAmps": (None, [_Q, _PR, _PR]),
    "initStateDebug": (None, [_Q]), "setAmps": (None, [_Q, _LL, _PR, _PR, _LL]), "setDensityAmps": (None, [_Q, _PR, _PR]),
    "initStateOfSingleQubit": (None, ["pQureg", _I, _I]), "compareStates": (_I, [_Q, _Q, _R]),
    # one-target gates
    "hadamard": (None, [_Q, _I]), "pauliX": (None, [_Q, _I]), "pauliY": (None, [_Q, _I]), "pauliZ": (None, [_Q, _I]),
    "sGate": (None, [_Q, _I]), "tGate": (None, [_Q, _I]), "phaseShift": (None, [_Q, _I, _R]),
    "rotateX": (None, [_Q, _I, _R]), "rotateY": (None, [_Q, _I, _R]), "rotateZ": (None, [_Q, _I, _R]),
    "rotateAroundAxis": (None, [_Q, _I, _R, _V]), "compactUnitary": (None, [_Q, _I, _CX, _CX]),
    "unitary": (None, [_Q, _I, _M2]),
    # controlled gates
    "controlledNot": (None, [_Q, _I, _I]), "controlledPauliY": (None, [_Q, _I, _I]),
    "controlledPhaseFlip": (None, [_Q, _I, _I]), "controlledPhaseShift": (None, [_Q, _I, _I, _R]),
    "controlledRotateX": (None, [_Q, _I, _I, _R]), "controlledRotateY": (None, [_Q, _I, _I, _R]),
    "controlledRotateZ": (None, [_Q, _I, _I, _R]), "controlledRotateAroundAxis": (None, [_Q, _I, _I, _R, _V]),
    "controlledCompactUnitary": (None, [_Q, _I, _I, _CX, _CX]), "controlledUnitary": (None, [_Q, _I, _I, _M2]),
    "multiControlledPhaseFlip": (None, [_Q, _PI, _I]), "multiControlledPhaseShift": (None, [_Q, _PI, _I, _R]),
    "multiControlledUnitary": (None, [_Q, _PI, _I, _I, _M2]),
    # noise
    "applyOneQubitDephaseError": (None, [_Q, _I, _R]), "applyOneQubitDepolariseError": (None, [_Q, _I, _R]),
    "applyOneQubitDampingError": (None, [_Q, _I, _R]), "applyTwoQubitDephaseError": (None, [_Q, _I, _I, _R]),
    "applyTwoQubitDepolariseError": (None, [_Q, _I, _I, _R]), "addDensityMatrix": (None, [_Q, _R, _Q]),
    # maths / measurement
    "calcFidelity": (_R, [_Q, _Q]), "calcInnerProduct": (_CX, [_Q, _Q]), "calcProbOfOutcome": (_R, [_Q, _I, _I]),
    "calcTotalProb": (_R, [_Q]), "calcPurity": (_R, [_Q]), "collapseToOutcome": (_R, [_Q, _I, _I]),
    "getAmp": (_CX, [_Q, _LL]), "getDensityAmp": (_CX, [_Q, _LL, _LL]), "getRealAmp": (_R, [_Q, _LL]),
    "getImagAmp": (_R, [_Q, _LL]), "getProbAmp": (_R, [_Q, _LL]), "measure": (_I, [_Q, _I]),
    "measureWithStats": (_I, [_Q, _I, _PR]),
    # RNG (exported by the reference and used by its seedQuEST known-answer test)
    "genrand_real1": ("double", []),
}

# entry points only quest-b200 libraries export (product and oracle); absent from the reference build
EXTENSIONS = {
    "mixDephasing": (None, [_Q, _I, _R]), "mixDepolarising": (None, [_Q, _I, _R]), "mixDamping": (None, [_Q, _I, _R]),
    "getEnvironmentString": (None, [_E, _Q, "cbuf"]),
    "qb200_copyAmpsToHost": (None, [_Q, _LL, _LL, _PR, _PR]),
}


class QuESTLibrary:
    """One loaded QuEST-compatible shared library plus typed call wrappers."""

    def __init__(self, path: str):
        if not os.path.isfile(path):
            raise FileNotFoundError(f"QuEST library not found: {path}")
        self.path = path
        # RTLD_LOCAL keeps the (identically named) symbols of several QuEST builds apart in one process
        self.lib = C.CDLL(path, mode=C.RTLD_LOCAL)
        prec_fn = self.lib.getQuEST_PREC
        prec_fn.restype = C.c_int
        self.prec = prec_fn()
        self.qreal = {1: C.c_float, 2: C.c_double}[self.prec]
        self.np_real = {1: np.float32, 2: np.float64}[self.prec]
        self.np_complex = {1: np.complex64, 2: np.complex128}[self.prec]
        self.eps = {1: 1e-5, 2: 1e-13}[self.prec]
        self.types = _make_types(self.qreal)
        for k, v in self.types.items():
            setattr(self, k, v)
        self._tokens = {
            "Qureg": self.Qureg, "QuESTEnv": self.QuESTEnv, "int": C.c_int, "ll": C.c_longlong, "qreal": self.qreal,
            "pqreal": C.POINTER(self.qreal), "pint": C.POINTER(C.c_int), "Complex": self.Complex,
            "ComplexMatrix2": self.ComplexMatrix2, "Vector": self.Vector, "pulong": C.POINTER(C.c_ulong),
            "cstr": C.c_char_p, "cbuf": C.c_char_p, "pQureg": C.POINTER(self.Qureg), "double": C.c_double,
            "voidp": C.c_void_p, "pll": C.POINTER(C.c_longlong), "pdouble": C.POINTER(C.c_double),
        }
        self.missing: list[str] = []
        self._declare(API, required=True)
        self._declare(EXTENSIONS, required=False)

    # -- declaration ---------------------------------------------------------------------------------
    def _resolve(self, tok):
        return None if tok is None else self._tokens[tok]

    def _declare(self, table, required):
        for name, (res, args) in table.items():
            try:
                fn = getattr(self.lib, name)
            except AttributeError:
                if required:
                    self.missing.append(name)
                continue
            fn.restype = self._resolve(res)
            fn.argtypes = [self._resolve(a) for a in args]
            setattr(self, name, fn)

    def declare(self, name: str, res, args: Sequence[str]):
        """Declare one more exported symbol (used for quest_b200.h extension entry points)."""
        fn = getattr(self.lib, name)
        fn.restype = self._resolve(res)
        fn.argtypes = [self._resolve(a) for a in args]
        setattr(self, name, fn)
        return fn

    def has(self, name: str) -> bool:
        return hasattr(self.lib, name)

    # -- argument builders ---------------------------------------------------------------------------
    def complex(self, z) -> "C.Structure":
        z = complex(z)
        return self.Complex(z.real, z.imag)

    def matrix(self, m) -> "C.Structure":
        m = np.asarray(m, dtype=complex).reshape(2, 2)
        return self.ComplexMatrix2(self.complex(m[0, 0]), self.complex(m[0, 1]), self.complex(m[1, 0]), self.complex(m[1, 1]))

    def vector(self, v) -> "C.Structure":
        return self.Vector(float(v[0]), float(v[1]), float(v[2]))

    def int_array(self, xs: Iterable[int]):
        xs = list(xs)
        return (C.c_int * len(xs))(*xs)

    def real_ptr(self, a: np.ndarray):
        assert a.dtype == self.np_real and a.flags.c_contiguous
        return a.ctypes.data_as(C.POINTER(self.qreal))

    # -- whole-state transfer --------------------------------------------------------------------------
    def set_state(self, qureg, amps: np.ndarray, start: int = 0):
        """Overwrite amplitudes [start, start+len) with a complex array (setAmps / setDensityAmps semantics)."""
        amps = np.asarray(amps)
        re = np.ascontiguousarray(amps.real, dtype=self.np_real)
        im = np.ascontiguousarray(amps.imag, dtype=self.np_real)
        if qureg.isDensityMatrix:
            assert start == 0 and len(re) == qureg.numAmpsTotal
            self.setDensityAmps(qureg, self.real_ptr(re), self.real_ptr(im))
        else:
            self.setAmps(qureg, start, self.real_ptr(re), self.real_ptr(im), len(re))

    def get_state(self, qureg) -> np.ndarray:
        """This rank's shard as a complex numpy array (density matrices: flattened column-major)."""
        n = int(qureg.numAmpsPerChunk)
        if hasattr(self, "qb200_copyAmpsToHost"):
            re = np.empty(n, dtype=self.np_real)
            im = np.empty(n, dtype=self.np_real)
            self.qb200_copyAmpsToHost(qureg, 0, n, self.real_ptr(re), self.real_ptr(im))
        else:  # the reference CPU build keeps the state in host arrays
            re = np.ctypeslib.as_array(qureg.stateVec.real, shape=(n,)).copy()
            im = np.ctypeslib.as_array(qureg.stateVec.imag, shape=(n,)).copy()
        return (re + 1j * im).astype(self.np_complex)

    def qasm_text(self, qureg) -> str:
        return qureg.qasmLog.contents.buffer.decode()


_cache: dict[str, QuESTLibrary] = {}


def load(path: str) -> QuESTLibrary:
    path = os.path.abspath(path)
    if path not in _cache:
        _cache[path] = QuESTLibrary(path)
    return _cache[path]


def load_product(prec: int = 2) -> QuESTLibrary:
    """The CUDA library.  Raises if it has not been built; there is no CPU stand-in."""
    path = PRODUCT_LIB if prec == 2 else PRODUCT_LIB_F32
    if prec == 2 and os.environ.get("QUEST_B200_LIB"):  # kernel A/B experiments (tools/ab_variants.sh): another build of the product
        path = os.environ["QUEST_B200_LIB"]
    if not os.path.isfile(path):
        raise RuntimeError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
            "quest-b200 has no CPU fallback.")
    return load(path)
