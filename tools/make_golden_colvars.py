"""Golden fixture for the SMD work observable: the first 201 rows (steps 0..10000) of the reference's
own NAMD/colvars record of replica 0, copied verbatim.  Data, not code; run here only (the reference tree does not
travel to the GPU box).

    python tools/make_golden_colvars.py
"""
import os

SRC = "/root/reference/All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/0/smd_out.colvars.traj"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden", "smd_out_head.colvars.traj")

if __name__ == "__main__":
    with open(SRC) as src, open(DST, "w") as dst:
        rows = 0
        for line in src:
            if line.startswith("#"):
                if rows == 0:
                    dst.write(line)
                continue
            dst.write(line)
            rows += 1
            if rows == 201:
                break
    print("wrote", DST, rows, "rows")
