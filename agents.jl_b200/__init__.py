"""agents.jl_b200 — host-side mirror of the Agents.jl interface for the
B200-native hot path (radius neighbour search + fused per-agent step).

Julia is not available in this image, so the host side above the C ABI
(include/agents_b200.h) is written in Python and mirrors the reference's
names, argument meaning and error behaviour for this path
(`ContinuousSpace`, `GridSpace`, `GridSpaceSingle`, `StandardABM`,
`nearby_ids`, `interacting_pairs`, `move_agent!`, `walk!`, `step!`).  The Julia
`ccall` glue a maintainer would add is in INTEGRATION.md.

Everything here calls the CUDA library through ctypes; nothing falls back to
the CPU and nothing imports oracle/.
"""
from ._lib import build, declared_symbols, lib  # noqa: F401
from .model import (  # noqa: F401
    AgentsError,
    ContinuousSpace,
    FlockRule,
    GridSpace,
    GridSpaceSingle,
    PairForceRule,
    SchellingRule,
    StandardABM,
    load_checkpoint,
)

__all__ = [
    "AgentsError", "ContinuousSpace", "GridSpace", "GridSpaceSingle", "StandardABM", "FlockRule",
    "PairForceRule", "SchellingRule", "load_checkpoint", "build", "lib", "declared_symbols",
]
