!> Stand-in for MOM_domains (MOM_tracer_advect.F90:10-11), one PE. Test infrastructure.
!!
!! create_group_pass records pointers to the fields; do_group_pass fills every halo point that has a
!! source in the global domain: cyclic wrap in x and/or y, tripolar north fold (i-reversal, sign change
!! of both components of a CGRID_NE vector pair), nothing for closed edges (the halo keeps its values,
!! as mpp_do_group_update leaves it).  Same rule as oracle/advect_ref.cpp locate().
module MOM_domains
implicit none ; private
public :: sum_across_PEs, max_across_PEs, create_group_pass, do_group_pass, group_pass_type, pass_var
public :: MOM_domain_type

integer, parameter :: MAXF = 256

type :: MOM_domain_type
  logical :: cyclic_x = .false., cyclic_y = .false., tripolar_n = .false., symmetric = .true.
  integer :: isc = 0, iec = 0, jsc = 0, jec = 0, isd = 0, ied = 0, jsd = 0, jed = 0
  integer :: IsdB = 0, JsdB = 0
end type MOM_domain_type

type :: p3d
  real, dimension(:,:,:), pointer :: p => NULL()
end type p3d

type :: group_pass_type
  integer :: ns = 0, nv = 0
  logical :: used = .false.   !< a pass has been done since the last create: the next create starts a new list
  type(p3d) :: s(MAXF), u(MAXF), v(MAXF)
end type group_pass_type

interface create_group_pass
  module procedure create_scalar_3d, create_vector_3d
end interface create_group_pass

interface pass_var
  module procedure pass_var_3d
end interface pass_var

interface sum_across_PEs
  module procedure sum_int_1d
end interface sum_across_PEs

interface max_across_PEs
  module procedure max_real_0d
end interface max_across_PEs

integer, parameter :: POS_H = 0, POS_U = 1, POS_V = 2

contains

subroutine restart_if_used(group)
  type(group_pass_type), intent(inout) :: group
  if (group%used) then
    group%ns = 0 ; group%nv = 0 ; group%used = .false.
  endif
end subroutine restart_if_used

subroutine create_scalar_3d(group, array, MOM_dom, sideflag, position, halo, clock)
  type(group_pass_type),  intent(inout) :: group
  real, dimension(:,:,:), target, intent(inout) :: array
  type(MOM_domain_type),  intent(inout) :: MOM_dom
  integer, optional,      intent(in)    :: sideflag, position, halo, clock
  call restart_if_used(group)
  if (group%ns >= MAXF) error stop "stub MOM_domains: too many scalar fields in one group"
  group%ns = group%ns + 1
  group%s(group%ns)%p => array
end subroutine create_scalar_3d

subroutine create_vector_3d(group, u_cmpt, v_cmpt, MOM_dom, direction, stagger, halo, clock)
  type(group_pass_type),  intent(inout) :: group
  real, dimension(:,:,:), target, intent(inout) :: u_cmpt, v_cmpt
  type(MOM_domain_type),  intent(inout) :: MOM_dom
  integer, optional,      intent(in)    :: direction, stagger, halo, clock
  call restart_if_used(group)
  if (group%nv >= MAXF) error stop "stub MOM_domains: too many vector fields in one group"
  group%nv = group%nv + 1
  group%u(group%nv)%p => u_cmpt
  group%v(group%nv)%p => v_cmpt
end subroutine create_vector_3d

subroutine do_group_pass(group, MOM_dom, clock)
  type(group_pass_type), intent(inout) :: group
  type(MOM_domain_type), intent(inout) :: MOM_dom
  integer, optional,     intent(in)    :: clock
  integer :: n
  do n=1,group%nv
    call fill(group%u(n)%p, MOM_dom, POS_U)
    call fill(group%v(n)%p, MOM_dom, POS_V)
  enddo
  do n=1,group%ns
    call fill(group%s(n)%p, MOM_dom, POS_H)
  enddo
  group%used = .true.
end subroutine do_group_pass

subroutine pass_var_3d(array, MOM_dom, sideflag, complete, position, halo, clock)
  real, dimension(:,:,:), intent(inout) :: array
  type(MOM_domain_type),  intent(inout) :: MOM_dom
  integer, optional,      intent(in)    :: sideflag, position, halo, clock
  logical, optional,      intent(in)    :: complete
  call fill(array, MOM_dom, POS_H)
end subroutine pass_var_3d

!> Fill the halo of one field.  a has lower bounds 1; local index i of the model is a-index i-ilo+1.
subroutine fill(a, D, pos)
  real, dimension(:,:,:), intent(inout) :: a
  type(MOM_domain_type),  intent(in)    :: D
  integer,                intent(in)    :: pos
  integer :: ilo, jlo, i, j, k, i0, j0, is, js, nh
  real :: sgn
  integer, allocatable :: di(:), dj(:), si(:), sj(:)
  real, allocatable :: sg(:), val(:)
  ilo = D%isd ; jlo = D%jsd
  if (pos == POS_U) ilo = D%IsdB
  if (pos == POS_V) jlo = D%JsdB
  allocate(di(size(a,1)*size(a,2)), dj(size(a,1)*size(a,2)), si(size(a,1)*size(a,2)), sj(size(a,1)*size(a,2)))
  allocate(sg(size(a,1)*size(a,2)), val(size(a,1)*size(a,2)))
  nh = 0
  do j=jlo,D%jed ; do i=ilo,D%ied
    i0 = D%isc ; j0 = D%jsc
    if (D%symmetric .and. pos == POS_U) i0 = D%isc-1
    if (D%symmetric .and. pos == POS_V) j0 = D%jsc-1
    if (i >= i0 .and. i <= D%iec .and. j >= j0 .and. j <= D%jec) cycle      ! not a halo point
    if (.not. locate(D, pos, i - (D%isc-1), j - (D%jsc-1), is, js, sgn)) cycle
    nh = nh + 1
    di(nh) = i-ilo+1 ; dj(nh) = j-jlo+1 ; si(nh) = is-ilo+1 ; sj(nh) = js-jlo+1 ; sg(nh) = sgn
  enddo ; enddo
  do k=1,size(a,3)
    do i=1,nh                      ! gather first, then write: sources are never halo points
      val(i) = a(si(i),sj(i),k)
      if (sg(i) < 0.0) val(i) = -val(i)
    enddo
    do i=1,nh
      a(di(i),dj(i),k) = val(i)
    enddo
  enddo
  deallocate(di, dj, si, sj, sg, val)
end subroutine fill

!> Owner (local indices is,js) of global point (ig,jg); .false. outside the global domain (closed edge).
logical function locate(D, pos, ig_in, jg_in, is, js, sgn)
  type(MOM_domain_type), intent(in)  :: D
  integer,               intent(in)  :: pos, ig_in, jg_in
  integer,               intent(out) :: is, js
  real,                  intent(out) :: sgn
  integer :: ig, jg, nig, njg, ilo, jlo, i0, j0
  ig = ig_in ; jg = jg_in ; sgn = 1.0 ; is = 0 ; js = 0 ; locate = .false.
  nig = D%iec - D%isc + 1 ; njg = D%jec - D%jsc + 1
  if (D%tripolar_n .and. jg > njg) then
    if (pos == POS_H) then
      ig = nig + 1 - ig ; jg = 2*njg + 1 - jg
    elseif (pos == POS_U) then
      ig = nig - ig ; jg = 2*njg + 1 - jg ; sgn = -1.0
    else
      ig = nig + 1 - ig ; jg = 2*njg - jg ; sgn = -1.0
    endif
  endif
  ilo = 1 ; jlo = 1
  if (pos == POS_U) ilo = 0
  if (pos == POS_V) jlo = 0
  if (D%cyclic_x) then
    do while (ig < 1) ; ig = ig + nig ; enddo
    do while (ig > nig) ; ig = ig - nig ; enddo
  elseif (ig < ilo .or. ig > nig) then
    return
  endif
  if (D%cyclic_y) then
    do while (jg < 1) ; jg = jg + njg ; enddo
    do while (jg > njg) ; jg = jg - njg ; enddo
  elseif (jg < jlo .or. jg > njg) then
    return
  endif
  is = ig + (D%isc-1) ; js = jg + (D%jsc-1)
  i0 = D%isc ; j0 = D%jsc
  if (pos == POS_U .and. ig == 0) i0 = D%isc-1      ! western domain-boundary face
  if (pos == POS_V .and. jg == 0) j0 = D%jsc-1      ! southern domain-boundary face
  locate = (is >= i0 .and. is <= D%iec .and. js >= j0 .and. js <= D%jec)
  ! a boundary face that exists only in symmetric memory
  if (locate .and. pos == POS_U .and. ig == 0 .and. .not.D%symmetric) locate = (D%isd <= D%isc-1)
  if (locate .and. pos == POS_V .and. jg == 0 .and. .not.D%symmetric) locate = (D%jsd <= D%jsc-1)
end function locate

subroutine sum_int_1d(field, length, pelist)
  integer, dimension(:), intent(inout) :: field
  integer,               intent(in)    :: length
  integer, optional,     intent(in)    :: pelist(:)
end subroutine sum_int_1d

subroutine max_real_0d(field, pelist)
  real,              intent(inout) :: field
  integer, optional, intent(in)    :: pelist(:)
end subroutine max_real_0d

end module MOM_domains
