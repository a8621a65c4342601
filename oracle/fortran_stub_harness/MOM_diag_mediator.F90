!> Stand-in for MOM_diag_mediator (MOM_tracer_advect.F90:8-9): only the two types appear in
!! declarations of the path; the procedures are imported but never called. Test infrastructure.
module MOM_diag_mediator
implicit none ; private
public :: post_data, query_averaging_enabled, diag_ctrl, register_diag_field, safe_alloc_ptr, time_type
type :: time_type
  integer :: seconds = 0, days = 0
end type time_type
type :: diag_ctrl
  integer :: unused = 0
end type diag_ctrl
contains
subroutine post_data(diag_field_id, field, diag)
  integer, intent(in) :: diag_field_id
  real, dimension(:,:,:), intent(in) :: field
  type(diag_ctrl), intent(in) :: diag
end subroutine post_data
logical function query_averaging_enabled(diag)
  type(diag_ctrl), intent(in) :: diag
  query_averaging_enabled = .false.
end function query_averaging_enabled
integer function register_diag_field(module_name, field_name)
  character(len=*), intent(in) :: module_name, field_name
  register_diag_field = -1
end function register_diag_field
subroutine safe_alloc_ptr(ptr, is, ie, js, je, nk)
  real, dimension(:,:,:), pointer :: ptr
  integer, intent(in) :: is, ie, js, je, nk
  if (.not.associated(ptr)) then
    allocate(ptr(is:ie,js:je,nk)) ; ptr(:,:,:) = 0.0
  endif
end subroutine safe_alloc_ptr
end module MOM_diag_mediator
