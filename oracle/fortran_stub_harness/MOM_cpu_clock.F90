!> Stand-in for MOM_cpu_clock (MOM_tracer_advect.F90:6-7). Test infrastructure.
module MOM_cpu_clock
implicit none ; private
public :: cpu_clock_id, cpu_clock_begin, cpu_clock_end, CLOCK_MODULE, CLOCK_ROUTINE
integer, parameter :: CLOCK_MODULE = 11, CLOCK_ROUTINE = 21
integer :: n_clocks = 0
contains
integer function cpu_clock_id(name, synchro_flag, grain)
  character(len=*),  intent(in) :: name
  integer, optional, intent(in) :: synchro_flag, grain
  n_clocks = n_clocks + 1
  cpu_clock_id = n_clocks
end function cpu_clock_id
subroutine cpu_clock_begin(id)
  integer, intent(in) :: id
end subroutine cpu_clock_begin
subroutine cpu_clock_end(id)
  integer, intent(in) :: id
end subroutine cpu_clock_end
end module MOM_cpu_clock
