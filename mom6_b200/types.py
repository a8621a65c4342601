"""Host-side mirrors of the MOM6 types that cross the advect_tracer boundary.

Only the members the hot path reads are kept:

* ``OceanGrid``      <- ocean_grid_type      (src/core/MOM_grid.F90:25-105)
* ``VerticalGrid``   <- verticalGrid_type    (src/core/MOM_verticalGrid.F90:212-215)
* ``MOMDomain``      <- MOM_domain_type      (src/framework/MOM_domains.F90:184-191)
* ``TracerType`` / ``TracerRegistry`` / ``register_tracer``
                     <- src/tracer/MOM_tracer_types.F90:9-130,
                        src/tracer/MOM_tracer_registry.F90:52-256

Array convention: every field is a C-ordered numpy array indexed ``[k, j, i]``
(2-D: ``[j, i]``), which is byte-for-byte the Fortran column-major array
``(i, j, k)`` the reference passes.  Integer bounds keep their Fortran
(1-based, inclusive) values, exactly as ``G%isc`` etc. hold them.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

MAX_FIELDS_ = 50  # config_src/memory/dynamic_symmetric/MOM_memory.h:21-23 (no limit on the GPU side)


@dataclass
class MOMDomain:
    """Topology and block layout of the horizontal decomposition."""
    nig: int
    njg: int
    layout: tuple = (1, 1)          # (pi, pj) == MOM6 LAYOUT
    cyclic_x: bool = False          # REENTRANT_X
    cyclic_y: bool = False          # REENTRANT_Y
    tripolar_n: bool = False        # TRIPOLAR_N
    symmetric: bool = True          # symmetric memory (SYMMETRIC_MEMORY_)
    pe: tuple = (0, 0)              # this rank's position in the layout


@dataclass
class VerticalGrid:
    ke: int
    Angstrom_H: float = 1.0e-10
    H_subroundoff: float = 1.0e-30

    @staticmethod
    def default(ke: int, Angstrom_H: float = 1.0e-10, m_to_H: float = 1.0) -> "VerticalGrid":
        # MOM_verticalGrid.F90:215  H_subroundoff = 1e-20 * max(Angstrom_H, m_to_H*1e-17)
        return VerticalGrid(ke, Angstrom_H, 1.0e-20 * max(Angstrom_H, m_to_H * 1.0e-17))


class OceanGrid:
    """The slice of ocean_grid_type that advect_tracer reads."""

    def __init__(self, ni: int, nj: int, halo: int = 4, symmetric: bool = True,
                 first_direction: int = 0, domain: Optional[MOMDomain] = None,
                 jhalo: Optional[int] = None):
        jhalo = halo if jhalo is None else jhalo
        # config_src/infra/FMS2/MOM_domain_infra.F90:1815-1821: isd = 1 locally
        self.isd, self.jsd = 1, 1
        self.isc, self.jsc = 1 + halo, 1 + jhalo
        self.iec, self.jec = self.isc + ni - 1, self.jsc + nj - 1
        self.ied, self.jed = self.iec + halo, self.jec + jhalo
        # src/framework/MOM_hor_index.F90:87-96
        self.symmetric = symmetric
        self.IsdB = self.isd - 1 if symmetric else self.isd
        self.JsdB = self.jsd - 1 if symmetric else self.jsd
        self.IedB, self.JedB = self.ied, self.jed
        self.first_direction = first_direction
        self.Domain = domain if domain is not None else MOMDomain(ni, nj, symmetric=symmetric)
        # global index = local index + off (single block: local isc <-> global 1)
        self.ioff = -(self.isc - 1)
        self.joff = -(self.jsc - 1)
        nid, njd = self.nid, self.njd
        self.areaT = np.ones((njd, nid))
        self.IareaT = np.ones((njd, nid))
        self.mask2dT = np.ones((njd, nid))
        self.mask2dCu = np.ones((njd, self.nIB))
        self.mask2dCv = np.ones((self.nJB, nid))

    # sizes of the data domain for each stagger
    @property
    def nid(self): return self.ied - self.isd + 1
    @property
    def njd(self): return self.jed - self.jsd + 1
    @property
    def nIB(self): return self.IedB - self.IsdB + 1
    @property
    def nJB(self): return self.JedB - self.JsdB + 1
    @property
    def ni(self): return self.iec - self.isc + 1
    @property
    def nj(self): return self.jec - self.jsc + 1

    def shape_h(self, nk=None): return (self.njd, self.nid) if nk is None else (nk, self.njd, self.nid)
    def shape_u(self, nk=None): return (self.njd, self.nIB) if nk is None else (nk, self.njd, self.nIB)
    def shape_v(self, nk=None): return (self.nJB, self.nid) if nk is None else (nk, self.nJB, self.nid)

    # python index of Fortran index
    def pi(self, i): return i - self.isd
    def pj(self, j): return j - self.jsd
    def pI(self, I): return I - self.IsdB
    def pJ(self, J): return J - self.JsdB

    def interior(self, a):
        """Computational-domain view of an h-point array."""
        return a[..., self.pj(self.jsc):self.pj(self.jec) + 1, self.pi(self.isc):self.pi(self.iec) + 1]

    def interior_u(self, a):
        """u faces isc-1..iec, rows jsc..jec."""
        return a[..., self.pj(self.jsc):self.pj(self.jec) + 1, self.pI(self.isc - 1):self.pI(self.iec) + 1]

    def interior_v(self, a):
        """v faces jsc-1..jec, columns isc..iec."""
        return a[..., self.pJ(self.jsc - 1):self.pJ(self.jec) + 1, self.pi(self.isc):self.pi(self.iec) + 1]


@dataclass
class TracerType:
    """tracer_type, MOM_tracer_types.F90:9-119 (hot-path members only)."""
    t: np.ndarray
    name: str = ""
    conc_underflow: float = 0.0
    ad_x: Optional[np.ndarray] = None
    ad_y: Optional[np.ndarray] = None
    ad2d_x: Optional[np.ndarray] = None
    ad2d_y: Optional[np.ndarray] = None
    advection_xy: Optional[np.ndarray] = None


@dataclass
class TracerRegistry:
    """tracer_registry_type, MOM_tracer_types.F90:122-130."""
    ntr: int = 0
    Tr: List[TracerType] = field(default_factory=list)
    locked: bool = False


def register_tracer(tr_ptr: np.ndarray, Reg: TracerRegistry, name: str = "",
                    underflow_conc: float = 0.0, max_fields: Optional[int] = None) -> TracerType:
    """register_tracer, MOM_tracer_registry.F90:52-256: the registry points at
    package-owned storage (``Tr%t => tr_ptr``, :232); nothing is copied."""
    if Reg.locked:
        raise RuntimeError("register_tracer was called for variable %s with a locked tracer registry." % name)
    if max_fields is not None and Reg.ntr >= max_fields:
        raise RuntimeError("Increase MAX_FIELDS_ in MOM_memory.h to at least %d" % (Reg.ntr + 1))
    if not (tr_ptr.flags["C_CONTIGUOUS"] and tr_ptr.dtype == np.float64):
        raise ValueError("tracer storage must be contiguous float64 [k, j, i]")
    tr = TracerType(t=tr_ptr, name=name or ("tr%d" % (Reg.ntr + 1)), conc_underflow=underflow_conc)
    Reg.Tr.append(tr)
    Reg.ntr += 1
    return tr


def lock_tracer_registry(Reg: TracerRegistry) -> None:
    """lock_tracer_registry, MOM_tracer_registry.F90:261."""
    Reg.locked = True


# ---- open boundaries (src/core/MOM_open_boundary.F90: the members advect_x / advect_y read) --------------
OBC_NONE, OBC_DIRECTION_N, OBC_DIRECTION_S, OBC_DIRECTION_E, OBC_DIRECTION_W = 0, 100, 200, 300, 400   # :79-83


@dataclass
class SegmentHI:
    """segment%HI (hor_index_type): the segment's own index ranges."""
    isd: int = 0
    ied: int = 0
    jsd: int = 0
    jed: int = 0
    IsdB: int = 0
    IedB: int = 0
    JsdB: int = 0
    JedB: int = 0


@dataclass
class OBCSegmentTracer:
    """OBC_segment_tracer_type, MOM_open_boundary.F90:119-133.  tres: [k, j, I] over (IsdB:IedB, jsd:jed, nk)
    for E/W segments, [k, J, i] over (isd:ied, JsdB:JedB, nk) for N/S segments; None == not allocated."""
    ntr_index: int = -1              # index in the global tracer registry, 1-based as in Fortran
    tres: Optional[np.ndarray] = None
    OBC_inflow_conc: float = 0.0


@dataclass
class OBCSegment:
    """OBC_segment_type, MOM_open_boundary.F90:145-260 (hot-path members)."""
    direction: int = OBC_NONE
    is_E_or_W: bool = False
    is_N_or_S: bool = False
    specified: bool = False
    HI: SegmentHI = field(default_factory=SegmentHI)
    tr_Reg: Optional[List[OBCSegmentTracer]] = None      # None == unassociated; ntseg = len(tr_Reg)


@dataclass
class OceanOBC:
    """ocean_OBC_type, MOM_open_boundary.F90:264-400 (hot-path members)."""
    OBC_pe: bool = True
    segment: List[OBCSegment] = field(default_factory=list)
    open_u_BCs_exist_globally: bool = False
    open_v_BCs_exist_globally: bool = False
    specified_u_BCs_exist_globally: bool = False
    specified_v_BCs_exist_globally: bool = False

    @property
    def number_of_segments(self) -> int:
        return len(self.segment)
