"""Known-answer rows of the reference's own tests for the force path (2 particles, tolerance decimal=4).

Source rows: /root/reference/src/pytest/test_pair.py:22-306, test_pair_aniso.py:21-110, test_bond.py:20-193.
Each pair row: (api class name, params, r_cut, shift, distance, energy, force) with the reference's convention:
particle 0 at -d/2, particle 1 at +d/2 on x; energies == [E/2, E/2]; forces == [[-F,0,0],[F,0,0]].
"""

PAIR_CASES = [
    # Colloid — test_pair.py:25-72
    ("Colloid", {"A": 100.0, "a_1": 0, "a_2": 0, "sigma": 2.0}, 6.0, False, 3.0, -0.2224, -0.4020),
    ("Colloid", {"A": 100.0, "a_1": 1.5, "a_2": 0, "sigma": 1.05}, 6.0, False, 3.0, -0.2757, -0.7107),
    ("Colloid", {"A": 100.0, "a_1": 0, "a_2": 1.5, "sigma": 1.05}, 6.0, False, 3.0, -0.2757, -0.7107),
    ("Colloid", {"A": 100.0, "a_1": 1.5, "a_2": 0.75, "sigma": 1.05}, 6.0, False, 3.0, -1.0366, -1.8267),
    ("Colloid", {"A": 100.0, "a_1": 1.5, "a_2": 0.75, "sigma": 1.05}, 6.0, False, 7.0, 0, 0),
    # DPDGeneralWeight (kT = 0, v = 0) — test_pair.py:78-107
    ("DPDGeneralWeight", {"A": 2.0, "gamma": 4.5, "s": 0.5}, 1.0, False, 0.5, 0.25, 1.0),
    ("DPDGeneralWeight", {"A": 25.0, "gamma": 4.5, "s": 2}, 1.0, False, 1.05, 0, 0),
    ("DPDGeneralWeight", {"A": 0.0, "gamma": 4.5, "s": 2}, 1.0, False, 0.5, 0, 0),
    # ExpandedYukawa — test_pair.py:113-171
    ("ExpandedYukawa", {"epsilon": 1.0, "kappa": 1.0, "delta": 1.0}, 3.0, False, 1.05, 19.024588490014263, 399.5163582902992),
    ("ExpandedYukawa", {"epsilon": 3.0, "kappa": 1.0, "delta": 1.0}, 3.0, False, 1.05, 57.07376547004279, 1198.5490748708976),
    ("ExpandedYukawa", {"epsilon": 1.0, "kappa": 1.0, "delta": 3.0}, 9.0, False, 3.05, 19.024588490014263, 399.5163582902992),
    ("ExpandedYukawa", {"epsilon": 1.0, "kappa": 3.0, "delta": 1.0}, 5.0, False, 1.05, 17.21415952850114, 395.9256691555259),
    ("ExpandedYukawa", {"epsilon": 1.0, "kappa": 1.0, "delta": 1.0}, 1.5, True, 1.05, 17.811527170588995, 399.5163583),
    ("ExpandedYukawa", {"epsilon": 1.0, "kappa": 1.0, "delta": 1.0}, 3.0, False, 4.0, 0, 0),
    # Hertz — test_pair.py:177-215
    ("Hertz", {"epsilon": 2.0}, 1.5, False, 1.05, 0.0985, 0.5477),
    ("Hertz", {"epsilon": 3.0}, 2.05, False, 1.05, 0.4985, 1.2464),
    ("Hertz", {"epsilon": 1.0}, 1.0, False, 1.05, 0, 0),
    ("Hertz", {"epsilon": 0.0}, 3.0, False, 1.05, 0, 0),
    # PerturbedLennardJones — test_pair.py:222-305
    ("PerturbedLennardJones", {"epsilon": 2.0, "sigma": 1.05, "attraction_scale_factor": 0.0}, 3.0, False, 1.05, 2.0, 45.7143),
    ("PerturbedLennardJones", {"epsilon": 2.0, "sigma": 1.05, "attraction_scale_factor": 0.5}, 3.0, False, 1.05, 1.0, 45.7143),
    ("PerturbedLennardJones", {"epsilon": 2.0, "sigma": 0.5, "attraction_scale_factor": 0.0}, 3.0, False, 1.05, 0, 0),
    ("PerturbedLennardJones", {"epsilon": 2.0, "sigma": 0.5, "attraction_scale_factor": 0.5}, 3.0, False, 1.05, -0.0460947, -0.260291),
    ("PerturbedLennardJones", {"epsilon": 2.0, "sigma": 1.05, "attraction_scale_factor": 0.5}, 3.0, True, 1.05, 1.00734, 45.7143),
    ("PerturbedLennardJones", {"epsilon": 2.0, "sigma": 0.85, "attraction_scale_factor": 0.5}, 3.0, True, 1.05, -0.806849, -2.81197),
    ("PerturbedLennardJones", {"epsilon": 1.0, "sigma": 1.0, "attraction_scale_factor": 0.5}, 1.0, False, 1.05, 0, 0),
    ("PerturbedLennardJones", {"epsilon": 0.0, "sigma": 1.0, "attraction_scale_factor": 0.5}, 3.0, False, 1.05, 0, 0),
]

# test_pair_aniso.py:21-110 — positions [-0.5,-0.10,-0.15], [0.5,0.10,0.15], both quaternions [1,0,0,0] (:123-125).
# Row: (params, r_cut, shift, energy, force on particle 1 or None, torque on both or None)
_TPM = {"M_d": 1.8341, "M_r": 0.0302, "r_eq": 1.0043, "omega": 5.0, "alpha": 0.40, "repulsion": False}
ANISO_POSITIONS = [[-0.5, -0.10, -0.15], [0.5, 0.10, 0.15]]
ANISO_CASES = [
    (dict(_TPM), 1.6, False, -0.20567 * 2, (-11.75766, -2.46991, -3.70487), (0.0, -0.08879, 0.05919)),
    (dict(_TPM), 1.10, True, -0.14195 * 2, None, None),
    (dict(_TPM), 1.0, True, 0.0, None, None),
    (dict(_TPM, M_d=0.0), 1.6, True, 0.0, None, None),
    (dict(_TPM, r_eq=1.1, omega=100.0), 1.6, False, -1.8341, (0.0, 0.0, 0.0), None),
]

# test_bond.py:20-193 — (api class, params, distance, energy, force); particle 1 (at +x) receives [F,0,0]
_Q = {"k": 1434.3, "r_0": 1.5, "b_1": -0.7589, "b_2": 0.0, "U_0": 67.2234, "sigma": 1.0, "epsilon": 1.0, "delta": 0.0}
BOND_CASES = [
    ("DoubleWell", {"r_0": 0.5, "r_1": 2.5, "U_1": 5.0, "U_tilt": 0.0}, 0.5, 0.0, 0.0),
    ("DoubleWell", {"r_0": 0.5, "r_1": 2.5, "U_1": 5.0, "U_tilt": 0.0}, 2.5, 5.0, 0.0),
    ("DoubleWell", {"r_0": 0.5, "r_1": 2.5, "U_1": 5.0, "U_tilt": 0.0}, 4.5, 0.0, 0.0),
    ("DoubleWell", {"r_0": 1.0, "r_1": 2.0, "U_1": 1.0, "U_tilt": 0.0}, 1.5, 0.5625, -1.5),
    ("DoubleWell", {"r_0": 1.0, "r_1": 2.0, "U_1": 1.0, "U_tilt": 0.0}, 2.5, 0.5625, 1.5),
    ("DoubleWell", {"r_0": 1.0, "r_1": 2.0, "U_1": 1.0, "U_tilt": 0.5}, 2.5, 1.03125, 0.25),
    ("Quartic", dict(_Q, sigma=0.0, epsilon=0.0), 1.0, 20.80586625, -99.2177025),
    ("Quartic", dict(_Q, k=0.0), 1.0, 68.2234, 24.0),
    ("Quartic", dict(_Q), 1.0, 21.80586625, -75.2177025),
    ("Quartic", dict(_Q, delta=0.5), 1.5, 21.80586625, -75.2177025),
    ("Quartic", dict(_Q), 1.5, 67.2234, 0.0),
    ("Quartic", dict(_Q), 1.5, 67.2234, 0.0),
    ("Quartic", dict(_Q, b_1=0.0, b_2=0.0), 1.25, 72.82613438, 89.64375),
]

# Random123 known-answer vectors for philox4x32_10 (kat_vectors of the Random123 distribution): (ctr, key, out)
PHILOX_KAT = [
    ((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
    ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
    ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0), (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
]
