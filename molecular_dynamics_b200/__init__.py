"""B200-native drop-in for the TorchMD force-field + integrator hot path of johncava/Molecular_Dynamics."""
