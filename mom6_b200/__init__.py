"""mom6_b200: B200-native MOM6 advect_tracer."""
from .types import (MOMDomain, OceanGrid, VerticalGrid, TracerType, TracerRegistry,
                    register_tracer, lock_tracer_registry)

__all__ = ["MOMDomain", "OceanGrid", "VerticalGrid", "TracerType", "TracerRegistry",
           "register_tracer", "lock_tracer_registry"]
