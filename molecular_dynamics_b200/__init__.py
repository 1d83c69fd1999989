"""B200-native drop-in for the TorchMD force-field + integrator hot path of johncava/Molecular_Dynamics.

    from molecular_dynamics_b200 import Forces, System, Integrator, maxwell_boltzmann, Parameters

mirrors ``torchmd.forces.Forces`` / ``torchmd.systems.System`` (the reference's TorchMD_Fix copies) and
upstream ``torchmd.integrator``.  The compute path is libb200md.so (hand-written sm_100a CUDA behind the
C-ABI of include/b200md.h); importing this package never imports the CPU oracle.
"""
from .energy import Energy
from .forces import Forces, SMDBias
from .integrator import BOLTZMAN, TIMEFACTOR, Integrator, kinetic_energy, kinetic_to_temp, maxwell_boltzmann
from .parameters import Parameters, parameters_from_prmtop, parameters_from_psf_yaml
from .minimizers import minimize_bfgs, minimize_fire
from .systems import System
from .sharding import gather_observables, jarzynski_free_energy, shard_range
from .utils import LogWriter, TrajectoryWriter, load_checkpoint, read_colvars_traj, save_checkpoint, write_colvars_traj
from .wrapper import Wrapper

__all__ = ["Forces", "SMDBias", "Energy", "System", "Integrator", "maxwell_boltzmann", "kinetic_energy", "kinetic_to_temp",
           "Parameters", "parameters_from_prmtop", "parameters_from_psf_yaml", "TIMEFACTOR", "BOLTZMAN", "minimize_bfgs",
           "minimize_fire", "LogWriter", "TrajectoryWriter", "save_checkpoint", "load_checkpoint", "Wrapper",
           "write_colvars_traj", "read_colvars_traj", "jarzynski_free_energy", "gather_observables", "shard_range"]
