!> Stand-in for MOM_error_handler (names imported at MOM_tracer_advect.F90:12). Test infrastructure.
module MOM_error_handler
implicit none ; private
public :: MOM_error, MOM_mesg, is_root_pe, FATAL, WARNING, NOTE
integer, parameter :: NOTE = 0, WARNING = 1, FATAL = 2
contains
subroutine MOM_error(level, message, all_print)
  integer,           intent(in) :: level
  character(len=*),  intent(in) :: message
  logical, optional, intent(in) :: all_print
  if (level == FATAL) then
    write(0,'("FATAL: ",a)') trim(message)
    error stop 1
  elseif (level == WARNING) then
    write(0,'("WARNING: ",a)') trim(message)
  else
    write(0,'("NOTE: ",a)') trim(message)
  endif
end subroutine MOM_error
subroutine MOM_mesg(message, verb, all_print)
  character(len=*),  intent(in) :: message
  integer, optional, intent(in) :: verb
  logical, optional, intent(in) :: all_print
  write(0,'(a)') trim(message)
end subroutine MOM_mesg
logical function is_root_pe()
  is_root_pe = .true.
end function is_root_pe
end module MOM_error_handler
