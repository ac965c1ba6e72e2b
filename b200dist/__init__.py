"""b200dist — a Blackwell-native (sm_100a) distributed-training framework.

Strategy coverage mirrors negin513/distributed-pytorch-hpc (DDP, FSDP, TP, SP, PP, hybrid 2-D
meshes); the execution layer is hand-written tcgen05/TMEM/TMA kernels plus peer-memory
collectives over NVLink 5 / NVSwitch.  See DESIGN.md.
"""
__version__ = "0.1.0"
