"""lammps-kokkos-port_b200: B200-native short-range MD hot path (libb200md.so + host mirror).

The directory name is not a valid Python identifier; import it through `b200pkg.load()` at
the repo root (registers it as `lammps_kokkos_port_b200`).
"""
from . import b200md, eam_host, inputs, molecule_host  # noqa: F401
