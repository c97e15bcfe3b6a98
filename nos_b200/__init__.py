"""nos_b200 -- B200-native partition-planning engine for nos (one hot path).

The product is libnosplan.so (nos_b200/csrc, C-ABI in include/nosplan.h).  The Python modules here are
host-side plumbing: ctypes bindings (capi), the logical->SoA packers (packing, quota_packing, victim_packing,
preempt_packing), synthetic snapshot generators (synth) and the multi-GPU drivers (sharded).
"""
__all__ = ["capi", "packing", "quota_packing", "victim_packing", "preempt_packing", "synth", "sharded", "geometries"]
