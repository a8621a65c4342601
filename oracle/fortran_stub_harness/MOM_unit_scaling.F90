!> Stand-in for MOM_unit_scaling (MOM_tracer_advect.F90:22). Test infrastructure.
module MOM_unit_scaling
implicit none ; private
public :: unit_scale_type
type :: unit_scale_type
  real :: s_to_T = 1.0
end type unit_scale_type
end module MOM_unit_scaling
