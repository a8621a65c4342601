!> Stand-in for MOM_file_parser (MOM_tracer_advect.F90:13): get_param for the three kinds
!! tracer_advect_init reads (:1113-1144), from a key table the driver fills. Test infrastructure.
module MOM_file_parser
use MOM_error_handler, only : MOM_error, FATAL
implicit none ; private
public :: get_param, log_version, param_file_type, set_param
integer, parameter :: MAXP = 32
type :: param_file_type
  integer :: n = 0
  character(len=64)  :: key(MAXP)
  character(len=128) :: val(MAXP)
end type param_file_type
interface get_param
  module procedure get_param_real, get_param_logical, get_param_char
end interface get_param
contains
subroutine set_param(pf, key, val)
  type(param_file_type), intent(inout) :: pf
  character(len=*),      intent(in)    :: key, val
  pf%n = pf%n + 1 ; pf%key(pf%n) = key ; pf%val(pf%n) = val
end subroutine set_param
integer function find(pf, key)
  type(param_file_type), intent(in) :: pf
  character(len=*),      intent(in) :: key
  integer :: n
  find = 0
  do n=1,pf%n ; if (trim(pf%key(n)) == trim(key)) find = n ; enddo
end function find
subroutine log_version(CS, modulename, version, desc, log_to_all, all_default, layout, debugging)
  type(param_file_type), intent(in) :: CS
  character(len=*),      intent(in) :: modulename, version
  character(len=*), optional, intent(in) :: desc
  logical, optional,     intent(in) :: log_to_all, all_default, layout, debugging
end subroutine log_version
subroutine get_param_real(CS, modulename, varname, value, desc, units, default, fail_if_missing, &
                          do_not_read, do_not_log, scale, unscaled, layoutParam, debuggingParam)
  type(param_file_type), intent(in)    :: CS
  character(len=*),      intent(in)    :: modulename, varname
  real,                  intent(inout) :: value
  character(len=*), optional, intent(in) :: desc, units
  real,    optional,     intent(in)    :: default, scale
  real,    optional,     intent(out)   :: unscaled
  logical, optional,     intent(in)    :: fail_if_missing, do_not_read, do_not_log, layoutParam, debuggingParam
  integer :: n
  n = find(CS, varname)
  if (n > 0) then
    read(CS%val(n), *) value
  elseif (present(default)) then
    value = default
  elseif (present(fail_if_missing)) then
    if (fail_if_missing) call MOM_error(FATAL, "get_param: "//trim(varname)//" is missing")
  endif
  if (present(unscaled)) unscaled = value
  if (present(scale)) value = scale*value
end subroutine get_param_real
subroutine get_param_logical(CS, modulename, varname, value, desc, units, default, fail_if_missing, &
                             do_not_read, do_not_log, layoutParam, debuggingParam)
  type(param_file_type), intent(in)    :: CS
  character(len=*),      intent(in)    :: modulename, varname
  logical,               intent(inout) :: value
  character(len=*), optional, intent(in) :: desc, units
  logical, optional,     intent(in)    :: default, fail_if_missing, do_not_read, do_not_log, layoutParam, debuggingParam
  integer :: n
  n = find(CS, varname)
  if (n > 0) then
    value = (index(CS%val(n), "T") > 0 .or. index(CS%val(n), "t") > 0)
  elseif (present(default)) then
    value = default
  endif
end subroutine get_param_logical
subroutine get_param_char(CS, modulename, varname, value, desc, units, default, fail_if_missing, &
                          do_not_read, do_not_log, layoutParam, debuggingParam)
  type(param_file_type), intent(in)    :: CS
  character(len=*),      intent(in)    :: modulename, varname
  character(len=*),      intent(inout) :: value
  character(len=*), optional, intent(in) :: desc, units, default
  logical, optional,     intent(in)    :: fail_if_missing, do_not_read, do_not_log, layoutParam, debuggingParam
  integer :: n
  n = find(CS, varname)
  if (n > 0) then
    value = trim(CS%val(n))
  elseif (present(default)) then
    value = default
  endif
end subroutine get_param_char
end module MOM_file_parser
