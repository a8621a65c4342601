!> Stand-in for MOM_open_boundary (MOM_tracer_advect.F90:15-17): the members read at
!! :442-478,580-628,825-862,965-1014. Test infrastructure.
module MOM_open_boundary
implicit none ; private
public :: ocean_OBC_type, OBC_segment_type, OBC_segment_tracer_type, segment_tracer_registry_type
public :: OBC_NONE, OBC_DIRECTION_N, OBC_DIRECTION_S, OBC_DIRECTION_E, OBC_DIRECTION_W
#include <MOM_memory.h>
integer, parameter :: OBC_NONE = 0
integer, parameter :: OBC_DIRECTION_N = 100, OBC_DIRECTION_S = 200, OBC_DIRECTION_E = 300, OBC_DIRECTION_W = 400
type :: hor_index_type
  integer :: isd, ied, jsd, jed, IsdB, IedB, JsdB, JedB
end type hor_index_type
type :: OBC_segment_tracer_type
  real, allocatable, dimension(:,:,:) :: tres
  real :: OBC_inflow_conc = 0.0
  integer :: ntr_index = -1
end type OBC_segment_tracer_type
type :: segment_tracer_registry_type
  integer :: ntseg = 0
  type(OBC_segment_tracer_type) :: Tr(MAX_FIELDS_)
end type segment_tracer_registry_type
type :: OBC_segment_type
  logical :: is_E_or_W = .false., is_N_or_S = .false., specified = .false., open = .false.
  integer :: direction = OBC_NONE
  type(hor_index_type) :: HI
  type(segment_tracer_registry_type), pointer :: tr_Reg => NULL()
end type OBC_segment_type
type :: ocean_OBC_type
  logical :: OBC_pe = .false.
  integer :: number_of_segments = 0
  logical :: open_u_BCs_exist_globally = .false., open_v_BCs_exist_globally = .false.
  logical :: specified_u_BCs_exist_globally = .false., specified_v_BCs_exist_globally = .false.
  type(OBC_segment_type), allocatable :: segment(:)
end type ocean_OBC_type
end module MOM_open_boundary
