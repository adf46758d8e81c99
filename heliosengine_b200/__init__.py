"""heliosengine_b200 — B200-native ECS hot path (archetype-table query iteration, deferred Commands,
double-buffered Messages, conflict-DAG scheduling) behind the HeliosEngine ecs API surface.

The product is heliosengine_b200/libhelios_gpu.so (hand-written sm_100a CUDA behind the C ABI in
include/helios_gpu.h); this package holds its ctypes bindings and the synthetic benchmark worlds.
"""
from . import capi  # noqa: F401
from .capi import World, Query, Schedule, load_gpu, fnv1a  # noqa: F401
