!> Stand-in for MOM_tracer_registry (MOM_tracer_advect.F90:21): the hot-path members of
!! tracer_type / tracer_registry_type (MOM_tracer_types.F90:9-130). Test infrastructure.
module MOM_tracer_registry
implicit none ; private
public :: tracer_registry_type, tracer_type
#include <MOM_memory.h>
type :: tracer_type
  real, dimension(:,:,:), pointer :: t => NULL()
  real, dimension(:,:,:), pointer :: ad_x => NULL(), ad_y => NULL(), advection_xy => NULL()
  real, dimension(:,:),   pointer :: ad2d_x => NULL(), ad2d_y => NULL()
  real :: conc_underflow = 0.0
end type tracer_type
type :: tracer_registry_type
  integer :: ntr = 0
  type(tracer_type) :: Tr(MAX_FIELDS_)
end type tracer_registry_type
end module MOM_tracer_registry
