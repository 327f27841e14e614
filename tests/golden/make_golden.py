#!/usr/bin/env python3
"""Convert the reference suite's golden vectors into committed fixtures, and record the reference's own
full-precision answers for the same cases.

Run HERE (the build container), where /root/reference and oracle/_ref/libQuEST.so exist:

    python tests/golden/make_golden.py

Inputs  : /root/reference/tests/unit/{state_vector,density_matrix}/{gates,maths,noise}/*.test
          (format: reference utilities/QuESTTest/QuESTCore.py:380-496 — header '# funcName', case count, then per
          case '<quregType>-<checks> <nQubits> [inline amps] args…' followed by the expected P / M / S blocks)
Outputs : tests/golden/ref_unit_tests.json.gz   every standard-format case: init state, typed arguments and the
                                                expected values exactly as printed in the .test file (13 dp)
          tests/golden/ref_unit_outputs.npz     for each case with an S block, the state the reference library
                                                produced here, as float64 (bit-level golden output)
          tests/golden/ref_circuits.npz         seeded random circuits (12-16 qubits) run through the reference
                                                library: final-state digests, sampled amplitudes, reductions
The Python-scripted reference tests (measure, measureWithStats, calcInnerProduct, calcFidelity, seedQuEST,
essential/*) are restated by hand in tests/test_reference_scripted.py.

Nothing on the GPU box reads /root/reference: tests load only the files written here.
"""
import gzip
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from quest_b200 import capi  # noqa: E402
from quest_b200.circuits import prepare, random_circuit, run_circuit, state_digest  # noqa: E402

REF_TESTS = "/root/reference/tests/unit"
STRIP = str.maketrans("", "", "[{()}]_|><")


def lines_of(path):
    """Non-empty lines with '#' comments removed (QuESTCore.py:187-205)."""
    out = []
    with open(path) as f:
        for raw in f:
            s = raw.split("#", 1)[0].strip()
            if s:
                out.append(s)
    return out


def func_name(path):
    with open(path) as f:
        for raw in f:
            s = raw.lstrip("# ").strip()
            if s:
                return s


def floats(tok):
    return [float(x) for x in tok.strip(",").split(",") if x]


def typed_args(func, toks):
    """Interpret the whitespace-split, bracket-stripped tokens with the C signature of `func`."""
    sig = capi.API[func][1][1:]  # drop the Qureg
    assert len(sig) == len(toks), (func, sig, toks)
    out = []
    for kind, tok in zip(sig, toks):
        if kind in ("int", "ll"):
            out.append(int(tok))
        elif kind == "qreal":
            out.append(float(tok))
        elif kind == "Complex":
            out.append({"complex": floats(tok)})
        elif kind == "ComplexMatrix2":
            out.append({"matrix": floats(tok)})
        elif kind == "Vector":
            out.append({"vector": floats(tok)})
        elif kind == "pint":
            out.append({"ints": [int(x) for x in tok.split(",")]})
        else:
            raise ValueError((func, kind))
    return out


def parse_file(path):
    func = func_name(path)
    if func.capitalize() == "Python" or func not in capi.API:
        return None
    ls = lines_of(path)
    n_cases = int(ls[0])
    pos = 1
    # the reference harness declares collapseToOutcome as returning nothing and checks P/M/S instead
    # (utilities/QuESTPy/QuESTFunc.py:99)
    ret = None if func == "collapseToOutcome" else capi.API[func][0]
    cases = []
    for _ in range(n_cases):
        toks = ls[pos].translate(STRIP).split()
        pos += 1
        head, nq, rest = toks[0], int(toks[1]), toks[2:]
        qtype, _, checks = head.partition("-")
        if nq == 0:
            continue  # the harness's "skip" marker (QuESTCore.py:394)
        dens = qtype.isupper()
        init = None
        if qtype in "CcBb":
            init, rest = rest[0], rest[1:]
            init = floats(init) if qtype in "Cc" else init
        case = {"func": func, "file": os.path.relpath(path, "/root/reference"), "qtype": qtype, "nq": nq,
                "density": dens, "init": init, "args": typed_args(func, rest)}
        n_amps = 4 ** nq if dens else 2 ** nq
        if ret is None:
            for chk in (checks or "S"):
                if chk in "Pp":
                    case["P"] = float(ls[pos]); pos += 1
                elif chk in "Mm":
                    case["M"] = [[float(x) for x in ls[pos + q].split()] for q in range(nq)]; pos += nq
                elif chk in "Ss":
                    vals = []
                    for a in range(n_amps):
                        vals += floats(ls[pos + a].translate(STRIP))
                    pos += n_amps
                    case["S"] = vals
        elif ret == "Complex":
            case["ret"] = floats(ls[pos].translate(STRIP)); pos += 1
        elif ret == "qreal":
            case["ret"] = float(ls[pos]); pos += 1
        elif ret == "int":
            case["ret"] = int(ls[pos]); pos += 1
        cases.append(case)
    return cases


def main():
    ref = capi.load(capi.REFERENCE_LIB)
    from tests.harness import run_golden_case  # the same runner the tests use

    all_cases = []
    for dirpath, _, files in sorted(os.walk(REF_TESTS)):
        for fn in sorted(files):
            if fn.endswith(".test"):
                parsed = parse_file(os.path.join(dirpath, fn))
                if parsed:
                    all_cases += parsed
    env = ref.createQuESTEnv()
    outputs = {}
    n_checks = 0
    for i, case in enumerate(all_cases):
        got = run_golden_case(ref, env, case)
        # the reference must reproduce its own golden file at the harness tolerance (1e-10, __main__.py:50)
        for key in ("P", "ret"):
            if key in case:
                assert np.allclose(got[key], case[key], atol=1e-10, rtol=0), (case["file"], key, got[key], case[key])
                n_checks += 1
        if "M" in case:
            assert np.allclose(got["M"], case["M"], atol=1e-10, rtol=0), case["file"]
            n_checks += 1
        if "S" in case:
            exp = np.array(case["S"][0::2]) + 1j * np.array(case["S"][1::2])
            assert np.allclose(got["S"], exp, atol=1e-10, rtol=0), case["file"]
            outputs[f"s{i}"] = got["S"]
            n_checks += 1
    with gzip.open(os.path.join(HERE, "ref_unit_tests.json.gz"), "wt") as f:
        json.dump(all_cases, f, separators=(",", ":"))
    np.savez_compressed(os.path.join(HERE, "ref_unit_outputs.npz"), **outputs)
    print(f"{len(all_cases)} cases, {n_checks} reference-verified checks -> ref_unit_tests.json.gz, ref_unit_outputs.npz")

    # seeded circuits through the reference library
    circ = {}
    for name, nq, dens, depth, seed in [("sv12", 12, False, 120, 11), ("sv16", 16, False, 160, 12),
                                        ("dm6", 6, True, 90, 13), ("dm8", 8, True, 120, 14)]:
        ops = random_circuit(nq, depth, seed, density=dens)
        q = ref.createDensityQureg(nq, env) if dens else ref.createQureg(nq, env)
        prepare(ref, q, seed)
        scal = run_circuit(ref, q, ops, seed=seed)
        st = ref.get_state(q)
        circ[name + "_state"] = st if st.size <= 1 << 16 else st[:: st.size >> 16]
        circ[name + "_digest"] = state_digest(st)
        circ[name + "_scalars"] = np.array(scal, dtype=np.float64)
        ref.destroyQureg(q, env)
    np.savez_compressed(os.path.join(HERE, "ref_circuits.npz"), **circ)
    print("circuits:", {k: v.shape for k, v in circ.items()})


if __name__ == "__main__":
    main()
