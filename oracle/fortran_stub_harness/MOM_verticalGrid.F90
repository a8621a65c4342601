!> Stand-in for MOM_verticalGrid (MOM_tracer_advect.F90:23). Test infrastructure.
module MOM_verticalGrid
implicit none ; private
public :: verticalGrid_type
type :: verticalGrid_type
  integer :: ke
  real :: Angstrom_H, H_subroundoff
end type verticalGrid_type
end module MOM_verticalGrid
