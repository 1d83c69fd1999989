"""Golden fixture for the batched torsion-angle output: 64 backbone frames of the
reference's own data set with the phi / psi angles stored beside the coordinates
(All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/processed_orient_phi_psi/0.npy, (1002, 40, 5) float32:
x, y, z in A, phi, psi in degrees per atom of the residue; produced by the reference's VMD / createPhiPsi tooling).
Copied verbatim (rows 0..63); data, not code.  Run here only.

    python tools/make_golden_phipsi.py
"""
import os

import numpy as np

SRC = "/root/reference/All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/processed_orient_phi_psi/0.npy"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "bb40_phi_psi.npy")

if __name__ == "__main__":
    a = np.load(SRC)
    np.save(DST, a[:64])
    print("wrote", DST, a[:64].shape, a.dtype)
