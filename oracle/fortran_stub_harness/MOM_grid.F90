!> Stand-in for MOM_grid (MOM_tracer_advect.F90:14): the members the path reads. Test infrastructure.
module MOM_grid
use MOM_domains, only : MOM_domain_type
implicit none ; private
public :: ocean_grid_type
type :: ocean_grid_type
  type(MOM_domain_type), pointer :: Domain => NULL()
  integer :: isc, iec, jsc, jec, isd, ied, jsd, jed
  integer :: IscB, IecB, JscB, JecB, IsdB, IedB, JsdB, JedB
  integer :: ke
  integer :: first_direction = 0
  real, allocatable, dimension(:,:) :: areaT, IareaT, mask2dT, mask2dCu, mask2dCv
end type ocean_grid_type
end module MOM_grid
