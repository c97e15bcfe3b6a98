"""nos_b200 -- B200-native partition-planning engine for nos (one hot path).

The product is libnosplan.so (nos_b200/csrc, C-ABI in include/nosplan.h).  The Python modules here are
host-side plumbing: ctypes bindings (capi), the logical->SoA packer (packing), synthetic snapshot
generators (synth) and a thin mirror of the reference's Planner interface (planner).
"""
__all__ = ["capi", "packing", "synth", "planner"]
