"""Fixtures for the CHAMBER prmtop route: the reference's ALA10 prmtop and the start structure of its
NAMD runs, copied verbatim (data, not code), plus NAMD's own step-0 energy line for that structure parsed out of the
run log.  Run here only (the reference tree does not travel to the GPU box).

    python tools/make_golden_chamber.py
"""
import json
import os
import shutil

REF = "/root/reference"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
FILES = {
    "full_da-1.3.prmtop": f"{REF}/V_Calculations/Test-6_full_system/data/full_da-1.3.prmtop",
    "smd_ini.pdb": f"{REF}/All_ML_Training_Data/210905_SMD_decaalanine/Build/smd_ini.pdb",
    # NAMD binary restart of replica 3 (rigidBonds ALL: every X-H bond sits at its equilibrium length)
    "smd_out.restart.coor": f"{REF}/All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/3/smd_out.restart.coor",
    "smd_out.restart.vel": f"{REF}/All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/3/smd_out.restart.vel",
    "smd_out.restart.xsc": f"{REF}/All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/3/smd_out.restart.xsc",
}
LOG = f"{REF}/All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/0/slurm_namdsim1-output.log"

if __name__ == "__main__":
    for name, src in FILES.items():
        shutil.copyfile(src, os.path.join(DST, name))
        os.chmod(os.path.join(DST, name), 0o644)
    titles = energy = None
    for line in open(LOG):
        if line.startswith("ETITLE:") and titles is None:
            titles = line.split()[1:]
        if line.startswith("ENERGY:") and energy is None:
            energy = [float(x) for x in line.split()[1:]]
    rec = {"source": LOG[len(REF) + 1:] + " (first ENERGY line: step 0 of the run that starts from my.pdb == smd_ini.pdb)",
           "note": "NAMD applies rigidBonds ALL to the hydrogens before this evaluation and uses GBIS + a switched VDW, "
                   "so BOND and ELECT are not comparable and the other columns agree only to a fraction of a percent",
           "energies": dict(zip(titles, energy))}
    # the production run of replica 3: its last ENERGY line (step 500000) belongs to the restart files copied above,
    # whose coordinates satisfy the constraints, so the bonded columns compare without the caveat of step 0
    log3 = f"{REF}/All_ML_Training_Data/210905_SMD_decaalanine/SMD/output/3/slurm_namdsim2-output.log"
    t3 = e3 = None
    for line in open(log3):
        if line.startswith("ETITLE:") and t3 is None:
            t3 = line.split()[1:]
        if line.startswith("ENERGY:") and line.split()[1] == "500000":
            e3 = [float(x) for x in line.split()[1:]]
    rec["restart_step_500000"] = {"source": log3[len(REF) + 1:] + " (ENERGY line of step 500000 = smd_out.restart.{coor,vel})",
                                  "energies": dict(zip(t3, e3))}
    with open(os.path.join(DST, "namd_step0_energies.json"), "w") as fh:
        json.dump(rec, fh, indent=1)
    print(rec["energies"])
