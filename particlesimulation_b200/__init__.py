"""particlesimulation_b200 -- B200-native per-timestep loop of `parsim`
(MiguelParece/ParticleSimulation, serial/parsim.cpp:422-448) behind a C ABI.

csrc/                hand-written sm_100a FP64 CUDA kernels + the extern "C" library + CLI
_lib.py              ctypes binding of libparsim_b200.so (fails loudly if it is not built)
simulation.py        host-side mirror of the reference's ParticleSimulation class
"""
from .simulation import ParticleSimulation, ParsimError, init_particles  # noqa: F401

__all__ = ["ParticleSimulation", "ParsimError", "init_particles"]
