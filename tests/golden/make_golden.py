"""Generates tests/golden/golden_v1.npz by running the REFERENCE's own CPU solver on the seeded cases.

Run in the build container only (needs oracle/_ref/libsart_ref.so, i.e. /root/reference at build time):

    make -C oracle && python tests/golden/make_golden.py

For each case in tests/cases.py it stores the reference's solution(s), status and iteration count, and
checksums of the regenerated inputs. The reference ships no golden vectors of its own,
so these outputs of its unmodified object code are the pin for both the C restatement and the CUDA engine.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from oracle.pyoracle import Reference, Restatement  # noqa: E402
from tests.cases import CASES, build_case, input_checksums  # noqa: E402


def main():
    ref, res = Reference(), Restatement()
    out = {}
    for name in CASES:
        case = build_case(name, res)
        x, status, iters, _ = ref.solve_frames(case["H"], case["frames"], x0=case["x0"], lap=case["lap"],
                                               no_guess=case["no_guess"], **case["kw"])
        out[name + "/x"] = x
        out[name + "/status"] = status
        out[name + "/iterations"] = iters
        out[name + "/checksums"] = input_checksums(case)
        print(f"{name:28s} status={status.tolist()} iterations={iters.tolist()} |x|={np.linalg.norm(x):.6g}")
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden_v1.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
