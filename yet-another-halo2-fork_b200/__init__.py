"""h2b200: B200-native MSM + NTT behind the halo2 `MsmAccel` / `EvaluationDomain` / `ParamsKZG` interfaces.

Importing this package does not touch the GPU; constructing a `Context` loads libh2b200.so and fails loudly
if it is missing or no sm_100 device is present (there is no CPU fallback).
"""
from . import evaluation, fields, synth
from . import plonk
from ._lib import H2bError, LIB_PATH, SYMBOLS, load
from .arithmetic import PolyOps
from .domain import EvaluationDomain
from .engine import (B200Engine, BaseDescriptor, CoeffsDescriptor, Context, PlonkEngine, PlonkEngineConfig, dist_all_gather_bytes,
                     shard_range, window_bits)
from .kzg import ParamsKZG

__all__ = ["evaluation", "fields", "synth", "plonk", "H2bError", "LIB_PATH", "SYMBOLS", "load", "EvaluationDomain", "B200Engine", "BaseDescriptor",
           "CoeffsDescriptor", "Context", "PlonkEngine", "PlonkEngineConfig", "ParamsKZG", "PolyOps", "shard_range", "window_bits", "dist_all_gather_bytes"]
