"""Writes tests/golden/ref_vectors.npz: outputs of the REFERENCE's own evaluator headers (oracle/_ref, compiled from
/root/reference/src by oracle/Makefile) on the seeded systems of tests/golden_systems.py. Run where /root/reference exists:

    make -C oracle ref && python tests/golden/make_golden.py

The fixture lets the C port and the CUDA path be checked against reference outputs on machines that have no
/root/reference and no prebuilt oracle/_ref (tests/test_golden_fixtures.py)."""

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

import golden_systems as GS  # noqa: E402
import oracle_api as O  # noqa: E402


def main():
    if not O.have_ref():
        raise SystemExit("oracle/_ref is not built (needs /root/reference): make -C oracle ref")
    out = {}
    for case in GS.all_cases():
        r = GS.run_oracle(case, "ref")
        s = GS.system(case)
        out[case["name"] + "/pos"] = s["pos"]  # guards against a silent change of the input recipe
        assert np.all(np.isfinite(r["force"])) and np.all(np.isfinite(r["virial"])), case["name"]
        out[case["name"] + "/force"] = np.asarray(r["force"], dtype=np.float64)
        out[case["name"] + "/virial"] = np.asarray(r["virial"], dtype=np.float64)
        if r["torque"] is not None:
            out[case["name"] + "/torque"] = np.asarray(r["torque"], dtype=np.float64)
    path = os.path.join(HERE, "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print(f"{len(GS.all_cases())} cases -> {path} ({os.path.getsize(path)} bytes)")


if __name__ == "__main__":
    main()
