"""gnn_uds_b200 - B200-native forward rollout of the GNN-UDS surrogate Emulator.

Only the hot path: ``Emulator`` construct/load/predict/simulate/rollout on hand-written
sm_100a CUDA behind a C ABI (include/gnnuds.h).  See DESIGN.md.
"""
__all__ = ["Emulator"]


def __getattr__(name):
    if name == "Emulator":
        from .emulator import Emulator
        return Emulator
    raise AttributeError(name)
