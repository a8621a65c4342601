!> Drop-in replacement of MOM6's module MOM_tracer_advect whose work is done on an
!! NVIDIA B200 through the C ABI of include/mom6adv.h (libmom6adv_b200.so).
!!
!! Same public names, dummy-argument lists and error behaviour as the reference
!! module (src/tracer/MOM_tracer_advect.F90:25-30, :52-86, :1090, :1153), so
!! step_MOM_tracer_dyn (src/core/MOM.F90:1438) and the offline tracer driver
!! (src/tracer/MOM_offline_main.F90:334,515,556,940,957) call it unchanged.
!! Build: list this file INSTEAD of src/tracer/MOM_tracer_advect.F90 (MOM6's
!! own same-named-module convention, config_src/external/README.md) and link
!! -lmom6adv_b200 -lcudart -lnccl.  See INTEGRATION.md.
!!
!! This file contains no arithmetic: it marshals G, GV, CS and the tracer
!! registry into plain-old-data descriptors and raw pointers.
module MOM_tracer_advect

use, intrinsic :: iso_c_binding, only : c_int, c_double, c_ptr, c_char, c_null_ptr, c_null_char, c_loc, &
                                        c_associated, c_f_pointer, c_long_long, c_intptr_t
use MOM_coms,            only : PE_here, root_PE, num_PEs, broadcast
use MOM_cpu_clock,       only : cpu_clock_id, cpu_clock_begin, cpu_clock_end
use MOM_cpu_clock,       only : CLOCK_MODULE, CLOCK_ROUTINE
use MOM_diag_mediator,   only : diag_ctrl, time_type
use MOM_domains,         only : sum_across_PEs
use MOM_error_handler,   only : MOM_error, FATAL, WARNING
use MOM_file_parser,     only : get_param, log_version, param_file_type
use MOM_grid,            only : ocean_grid_type
use MOM_open_boundary,   only : ocean_OBC_type, OBC_segment_type
use MOM_tracer_registry, only : tracer_registry_type, tracer_type
use MOM_unit_scaling,    only : unit_scale_type
use MOM_verticalGrid,    only : verticalGrid_type
implicit none ; private

#include <MOM_memory.h>

public advect_tracer
public tracer_advect_init
public tracer_advect_end
public tracer_hordiff_iterate
public offline_advection_iterate

!> Control structure: the reference's members (:30-40) plus the GPU plan handle.
type, public :: tracer_advect_CS ; private
  real    :: dt                      !< The baroclinic dynamics time step [T ~> s]
  type(diag_ctrl), pointer :: diag => NULL() !< Diagnostics control structure
  logical :: debug                   !< If true, write verbose checksums
  logical :: usePPM                  !< If true, use PPM instead of PLM
  logical :: useHuynh                !< If true, use the Huynh scheme for PPM interface values
  logical :: useHuynhStencilBug = .false. !< Legacy too-narrow stencil for PPM:H3
  type(c_ptr) :: plan = c_null_ptr   !< mom6adv_plan* (created on the first advect_tracer call)
  integer :: ntr_bound = -1          !< Reg%ntr the plan was last bound with
  integer(kind=c_intptr_t) :: reg_hash = 0 !< signature of the registry (pointers, underflow values) at the last bind
  logical :: obc_was_set = .false.   !< an OBC table is active in the plan
  logical :: reentrant_x = .false.   !< REENTRANT_X, as MOM_domains_init reads it (src/framework/MOM_domains.F90:184-191)
  logical :: reentrant_y = .false.   !< REENTRANT_Y
  logical :: tripolar_N = .false.    !< TRIPOLAR_N
end type tracer_advect_CS

integer :: id_clock_advect !< (Ocean advect tracer): the whole GPU call incl. host<->device copies
integer :: id_clock_pass   !< (Ocean tracer halo updates): kept for clock-table compatibility
integer :: id_clock_sync   !< (Ocean tracer global synch): kept for clock-table compatibility

! ---- POD descriptors of include/mom6adv.h ---------------------------------------------
type, bind(c) :: mom6adv_grid
  integer(c_int) :: isc, iec, jsc, jec, isd, ied, jsd, jed, IsdB, IedB, JsdB, JedB, nk
  type(c_ptr)    :: areaT, IareaT, mask2dT, mask2dCu, mask2dCv
  real(c_double) :: Angstrom_H, H_subroundoff
end type mom6adv_grid

type, bind(c) :: mom6adv_config
  real(c_double) :: dt_baroclinic
  integer(c_int) :: usePPM, useHuynh, useHuynhStencilBug, debug, first_direction
end type mom6adv_config

type, bind(c) :: mom6adv_domain
  integer(c_int) :: nig, njg, npi, npj, pi, pj, ioff, joff, cyclic_x, cyclic_y, tripolar_n
end type mom6adv_domain
type, bind(c) :: mom6adv_obc_tracer
  integer(c_int) :: ntr_index
  type(c_ptr)    :: tres
  real(c_double) :: OBC_inflow_conc
end type mom6adv_obc_tracer
type, bind(c) :: mom6adv_obc_segment
  integer(c_int) :: is_E_or_W, is_N_or_S, direction, specified
  integer(c_int) :: isd, ied, jsd, jed, IsdB, IedB, JsdB, JedB
  integer(c_int) :: ntseg
  type(c_ptr)    :: Tr
end type mom6adv_obc_segment
type, bind(c) :: mom6adv_obc
  integer(c_int) :: OBC_pe, number_of_segments
  integer(c_int) :: open_u_BCs_exist_globally, open_v_BCs_exist_globally
  integer(c_int) :: specified_u_BCs_exist_globally, specified_v_BCs_exist_globally
  type(c_ptr)    :: segment
end type mom6adv_obc

! ---- the C entry points (style of src/framework/posix.F90:52-229) ----------------------
interface
  function mom6adv_plan_create(g, c, dom, device, plan) bind(c, name="mom6adv_plan_create") result(rc)
    import :: c_int, c_ptr, mom6adv_grid, mom6adv_config, mom6adv_domain
    type(mom6adv_grid),   intent(in)  :: g
    type(mom6adv_config), intent(in)  :: c
    type(mom6adv_domain), intent(in)  :: dom
    integer(c_int), value             :: device
    type(c_ptr),          intent(out) :: plan
    integer(c_int) :: rc
  end function mom6adv_plan_create

  function mom6adv_plan_destroy(plan) bind(c, name="mom6adv_plan_destroy") result(rc)
    import :: c_int, c_ptr
    type(c_ptr), value :: plan
    integer(c_int) :: rc
  end function mom6adv_plan_destroy

  function mom6adv_last_error(plan) bind(c, name="mom6adv_last_error") result(msg)
    import :: c_ptr
    type(c_ptr), value :: plan
    type(c_ptr) :: msg
  end function mom6adv_last_error

  function mom6adv_comm_unique_id(id128) bind(c, name="mom6adv_comm_unique_id") result(rc)
    import :: c_int, c_char
    character(kind=c_char), intent(out) :: id128(128)
    integer(c_int) :: rc
  end function mom6adv_comm_unique_id

  function mom6adv_comm_init(plan, id128, ibeg, iend, jbeg, jend) bind(c, name="mom6adv_comm_init") result(rc)
    import :: c_int, c_char, c_ptr
    type(c_ptr), value :: plan
    character(kind=c_char), intent(in) :: id128(128)
    integer(c_int), intent(in) :: ibeg(*), iend(*), jbeg(*), jend(*)
    integer(c_int) :: rc
  end function mom6adv_comm_init

  function mom6adv_set_obc(plan, obc) bind(c, name="mom6adv_set_obc") result(rc)
    import :: c_int, c_ptr
    type(c_ptr), value :: plan
    type(c_ptr), value :: obc           !< c_loc of a mom6adv_obc, or c_null_ptr (unassociated OBC)
    integer(c_int) :: rc
  end function mom6adv_set_obc

  function mom6adv_set_ntr(plan, ntr) bind(c, name="mom6adv_set_ntr") result(rc)
    import :: c_int, c_ptr
    type(c_ptr), value :: plan
    integer(c_int), value :: ntr
    integer(c_int) :: rc
  end function mom6adv_set_ntr

  function mom6adv_bind_tracer(plan, m, t, conc_underflow, ad_x, ad_y, ad2d_x, ad2d_y, advection_xy) &
           bind(c, name="mom6adv_bind_tracer") result(rc)
    import :: c_int, c_ptr, c_double
    type(c_ptr), value :: plan
    integer(c_int), value :: m          !< 0-based tracer index
    type(c_ptr), value :: t, ad_x, ad_y, ad2d_x, ad2d_y, advection_xy   !< c_null_ptr == unassociated
    real(c_double), value :: conc_underflow
    integer(c_int) :: rc
  end function mom6adv_bind_tracer

  function mom6adv_advect(plan, h_end, uhtr, vhtr, dt, x_first, vol_prev, max_iter_in, update_vol_prev, &
                          uhr_out, vhr_out, n_iter) bind(c, name="mom6adv_advect") result(rc)
    import :: c_int, c_ptr, c_double
    type(c_ptr), value :: plan, h_end, uhtr, vhtr, vol_prev, uhr_out, vhr_out
    real(c_double), value :: dt
    integer(c_int), value :: x_first          !< -1 absent, 0 / 1
    integer(c_int), value :: max_iter_in      !< < 0 absent
    integer(c_int), value :: update_vol_prev  !< -1 absent, 0 / 1
    integer(c_int), intent(out) :: n_iter
    integer(c_int) :: rc
  end function mom6adv_advect

  !> uhtr += uh*dt ; vhtr += vh*dt on the resident fields (replaces MOM_dynamics_split_RK2.F90:1045-1052)
  function mom6adv_accumulate_transports(plan, uh, vh, dt) bind(c, name="mom6adv_accumulate_transports") result(rc)
    import :: c_int, c_ptr, c_double
    type(c_ptr), value :: plan, uh, vh
    real(c_double), value :: dt
    integer(c_int) :: rc
  end function mom6adv_accumulate_transports

  !> CS%uhtr(:,:,:) = 0.0 ; CS%vhtr(:,:,:) = 0.0 on the resident fields (replaces MOM.F90:1465-1466)
  function mom6adv_zero_transports(plan) bind(c, name="mom6adv_zero_transports") result(rc)
    import :: c_int, c_ptr
    type(c_ptr), value :: plan
    integer(c_int) :: rc
  end function mom6adv_zero_transports

  !> work2d of id_adv_xy_2d (kind 0), id_zint (1), id_zint_100m (2) for resident tracer m (0-based)
  !! (replaces the vertical sums of post_tracer_transport_diagnostics, MOM_tracer_registry.F90:745-811)
  function mom6adv_column_diag(plan, m, kind, h_diag, H_to_m, out2d) bind(c, name="mom6adv_column_diag") result(rc)
    import :: c_int, c_ptr, c_double
    type(c_ptr), value :: plan, h_diag, out2d
    integer(c_int), value :: m, kind
    real(c_double), value :: H_to_m
    integer(c_int) :: rc
  end function mom6adv_column_diag

  !> The diffusion iteration of tracer_hordiff on the resident tracers (replaces MOM_tracer_hor_diff.F90:539-607)
  function mom6adv_hordiff(plan, h, khdt_x, khdt_y, num_itts, h_neglect, diffuse_ml_interior, nkml, nk_rho_varies, &
                           ml_khtr_scale, want_df) bind(c, name="mom6adv_hordiff") result(rc)
    import :: c_int, c_ptr, c_double
    type(c_ptr), value :: plan, h, khdt_x, khdt_y   !< h = C_NULL_PTR: the resident h_end of the last advect_tracer call
    integer(c_int), value :: num_itts, diffuse_ml_interior, nkml, nk_rho_varies, want_df
    real(c_double), value :: h_neglect, ml_khtr_scale
    integer(c_int) :: rc
  end function mom6adv_hordiff

  !> The iteration loop of offline_advection_ale on the resident fields (replaces MOM_offline_main.F90:320-411)
  function mom6adv_offline_advect(plan, dt_offline, num_off_iter, off_ale_mod, x_before_y, min_residual, h_to_kg_m2, &
                                  l_to_m, kg_m2_to_h, m_to_l, converged, n_iter, tot_residual) &
                                  bind(c, name="mom6adv_offline_advect") result(rc)
    import :: c_int, c_ptr, c_double
    type(c_ptr), value :: plan, tot_residual    !< tot_residual: c_loc of num_off_iter+1 reals, or c_null_ptr
    real(c_double), value :: dt_offline, min_residual, h_to_kg_m2, l_to_m, kg_m2_to_h, m_to_l
    integer(c_int), value :: num_off_iter, off_ale_mod, x_before_y
    integer(c_int), intent(out) :: converged, n_iter
    integer(c_int) :: rc
  end function mom6adv_offline_advect

  function mom6adv_upload_field(plan, field, src) bind(c, name="mom6adv_upload_field") result(rc)
    import :: c_int, c_ptr
    type(c_ptr), value :: plan, src ; integer(c_int), value :: field ; integer(c_int) :: rc
  end function mom6adv_upload_field
  function mom6adv_download_field(plan, field, dst) bind(c, name="mom6adv_download_field") result(rc)
    import :: c_int, c_ptr
    type(c_ptr), value :: plan, dst ; integer(c_int), value :: field ; integer(c_int) :: rc
  end function mom6adv_download_field
end interface

contains

!> Raise MOM_error(FATAL) with the library's message when a C call failed.
subroutine check_rc(rc, plan, where)
  integer(c_int),   intent(in) :: rc    !< status returned by the C ABI (0 == OK)
  type(c_ptr),      intent(in) :: plan  !< plan handle (may be null)
  character(len=*), intent(in) :: where !< name of the caller
  character(kind=c_char), pointer :: cmsg(:)
  character(len=512) :: msg
  type(c_ptr) :: p
  integer :: n
  if (rc == 0) return
  msg = "" ; p = mom6adv_last_error(plan)
  if (c_associated(p)) then
    call c_f_pointer(p, cmsg, [512])
    do n=1,512 ; if (cmsg(n) == c_null_char) exit ; msg(n:n) = cmsg(n) ; enddo
  endif
  call MOM_error(FATAL, "MOM_tracer_advect (B200), "//trim(where)//": "//trim(msg))
end subroutine check_rc

!> Create the GPU plan: grid descriptors, HBM mirrors and, with more than one PE,
!! the halo-exchange lists and the NCCL communicator (stands in for the
!! create_group_pass calls of the reference, :140-144).
subroutine make_plan(G, GV, CS)
  type(ocean_grid_type), target, intent(inout) :: G   !< ocean grid structure
  type(verticalGrid_type),       intent(in)    :: GV  !< ocean vertical grid structure
  type(tracer_advect_CS),        intent(inout) :: CS  !< control structure
  type(mom6adv_grid)   :: cg
  type(mom6adv_config) :: cc
  type(mom6adv_domain) :: cd
  integer, allocatable :: ext(:)
  integer(c_int), allocatable :: ibeg(:), iend(:), jbeg(:), jend(:)
  character(kind=c_char) :: id(128)
  integer :: idw(32), npi, npj, rank, pi, pj, n
  cg%isc = G%isc ; cg%iec = G%iec ; cg%jsc = G%jsc ; cg%jec = G%jec
  cg%isd = G%isd ; cg%ied = G%ied ; cg%jsd = G%jsd ; cg%jed = G%jed
  cg%IsdB = G%IsdB ; cg%IedB = G%IedB ; cg%JsdB = G%JsdB ; cg%JedB = G%JedB
  cg%nk = GV%ke
  cg%areaT = c_loc(G%areaT) ; cg%IareaT = c_loc(G%IareaT) ; cg%mask2dT = c_loc(G%mask2dT)
  cg%mask2dCu = c_loc(G%mask2dCu) ; cg%mask2dCv = c_loc(G%mask2dCv)
  cg%Angstrom_H = GV%Angstrom_H ; cg%H_subroundoff = GV%H_subroundoff
  cc%dt_baroclinic = CS%dt
  cc%usePPM = merge(1, 0, CS%usePPM) ; cc%useHuynh = merge(1, 0, CS%useHuynh)
  cc%useHuynhStencilBug = merge(1, 0, CS%useHuynhStencilBug) ; cc%debug = merge(1, 0, CS%debug)
  cc%first_direction = G%first_direction
  npi = G%Domain%layout(1) ; npj = G%Domain%layout(2)
  rank = PE_here() - root_PE()              ! FMS lays PEs out with the i-position varying fastest
  pi = mod(rank, npi) ; pj = rank / npi
  cd%nig = G%Domain%niglobal ; cd%njg = G%Domain%njglobal
  cd%npi = npi ; cd%npj = npj ; cd%pi = pi ; cd%pj = pj
  cd%ioff = G%idg_offset ; cd%joff = G%jdg_offset        ! global index = local index + offset
  ! topology: the parameters create_MOM_domain turns into X_FLAGS / Y_FLAGS (config_src/infra/FMS2/MOM_domain_infra.F90:
  ! 1345-1366), read by tracer_advect_init from the parameter file -- MOM6 keeps FMS behind config_src/infra, so no
  ! mpp_domains_mod constant is used here
  cd%cyclic_x = merge(1, 0, CS%reentrant_x)
  cd%cyclic_y = merge(1, 0, CS%reentrant_y .and. .not.CS%tripolar_N)
  cd%tripolar_n = merge(1, 0, CS%tripolar_N)
  call check_rc(mom6adv_plan_create(cg, cc, cd, -1_c_int, CS%plan), c_null_ptr, "plan_create")
  if (npi*npj > 1) then
    ! global compute extents of every layout column / row: each PE contributes its own, the rest are zero
    allocate(ext(2*(npi+npj))) ; ext(:) = 0
    if (pj == 0) then ; ext(pi+1) = G%isc + G%idg_offset ; ext(npi+pi+1) = G%iec + G%idg_offset ; endif
    if (pi == 0) then ; ext(2*npi+pj+1) = G%jsc + G%jdg_offset ; ext(2*npi+npj+pj+1) = G%jec + G%jdg_offset ; endif
    call sum_across_PEs(ext, 2*(npi+npj))
    allocate(ibeg(npi), iend(npi), jbeg(npj), jend(npj))
    ibeg(:) = ext(1:npi) ; iend(:) = ext(npi+1:2*npi)
    jbeg(:) = ext(2*npi+1:2*npi+npj) ; jend(:) = ext(2*npi+npj+1:2*(npi+npj))
    ! the NCCL id: drawn on the root PE, broadcast as 32 default integers
    if (rank == 0) call check_rc(mom6adv_comm_unique_id(id), CS%plan, "comm_unique_id")
    idw = transfer(id, idw)
    call broadcast(idw, 32, root_PE())
    id = transfer(idw, id)
    call check_rc(mom6adv_comm_init(CS%plan, id, ibeg, iend, jbeg, jend), CS%plan, "comm_init")
  endif
end subroutine make_plan

!> A cheap signature of the registry: the plan is re-bound only when ntr, a tracer's storage, its underflow value or the
!! address of any of its diagnostic arrays changed.  Circular shifts and exclusive-ors only: no signed overflow.
function registry_hash(Reg) result(h)
  type(tracer_registry_type), intent(in) :: Reg !< tracer registry
  integer(kind=c_intptr_t) :: h
  integer :: m
  h = int(Reg%ntr, c_intptr_t)
  do m=1,Reg%ntr
    call mix(h, transfer(c_loc(Reg%Tr(m)%t), h))
    call mix(h, transfer(real(Reg%Tr(m)%conc_underflow, c_double), h))
    call mix(h, transfer(ptr3(Reg%Tr(m)%ad_x), h)) ; call mix(h, transfer(ptr3(Reg%Tr(m)%ad_y), h))
    call mix(h, transfer(ptr2(Reg%Tr(m)%ad2d_x), h)) ; call mix(h, transfer(ptr2(Reg%Tr(m)%ad2d_y), h))
    call mix(h, transfer(ptr3(Reg%Tr(m)%advection_xy), h))
  enddo
contains
  subroutine mix(acc, v)
    integer(kind=c_intptr_t), intent(inout) :: acc !< running signature
    integer(kind=c_intptr_t), intent(in)    :: v   !< value folded in
    acc = ieor(ishftc(acc, 7), v)
  end subroutine mix
end function registry_hash

!> Pointer of an optional diagnostic array, null when unassociated.
function ptr3(a) result(p)
  real, dimension(:,:,:), pointer :: a !< diagnostic array or null
  type(c_ptr) :: p
  p = c_null_ptr ; if (associated(a)) p = c_loc(a)
end function ptr3
function ptr2(a) result(p)
  real, dimension(:,:), pointer :: a !< diagnostic array or null
  type(c_ptr) :: p
  p = c_null_ptr ; if (associated(a)) p = c_loc(a)
end function ptr2

!> Horizontal tracer advection on the GPU; argument list of the reference (:52-86).
subroutine advect_tracer(h_end, uhtr, vhtr, OBC, dt, G, GV, US, CS, Reg, x_first_in, &
                         vol_prev, max_iter_in, update_vol_prev, uhr_out, vhr_out)
  type(ocean_grid_type), target, intent(inout) :: G     !< ocean grid structure
  type(verticalGrid_type), intent(in)    :: GV    !< ocean vertical grid structure
  real, dimension(SZI_(G),SZJ_(G),SZK_(GV)), target, &
                           intent(in)    :: h_end !< layer thickness after advection [H ~> m or kg m-2]
  real, dimension(SZIB_(G),SZJ_(G),SZK_(GV)), target, &
                           intent(in)    :: uhtr  !< accumulated zonal volume/mass flux [H L2 ~> m3 or kg]
  real, dimension(SZI_(G),SZJB_(G),SZK_(GV)), target, &
                           intent(in)    :: vhtr  !< accumulated meridional volume/mass flux [H L2 ~> m3 or kg]
  type(ocean_OBC_type),    pointer       :: OBC   !< open boundary condition data
  real,                    intent(in)    :: dt    !< time increment [T ~> s]
  type(unit_scale_type),   intent(in)    :: US    !< dimensional unit scaling type
  type(tracer_advect_CS),  pointer       :: CS    !< control structure for this module
  type(tracer_registry_type), pointer    :: Reg   !< tracer registry
  logical,       optional, intent(in)    :: x_first_in !< override of the first direction
  real, dimension(SZI_(G),SZJ_(G),SZK_(GV)), target, &
                 optional, intent(inout) :: vol_prev !< cell volume before advection [H L2 ~> m3 or kg]
  integer,       optional, intent(in)    :: max_iter_in !< maximum number of advection iterations
  logical,       optional, intent(in)    :: update_vol_prev !< if present and true, update vol_prev
  real, dimension(SZIB_(G),SZJ_(G),SZK_(GV)), target, &
                 optional, intent(out)   :: uhr_out !< remaining zonal flux [H L2 ~> m3 or kg]
  real, dimension(SZI_(G),SZJB_(G),SZK_(GV)), target, &
                 optional, intent(out)   :: vhr_out !< remaining meridional flux [H L2 ~> m3 or kg]
  integer(c_int) :: xf, mi, uv, nit
  integer(kind=c_intptr_t) :: h
  type(c_ptr) :: p_vol, p_uo, p_vo
  integer :: m

  if (.not. associated(CS)) call MOM_error(FATAL, "MOM_tracer_advect: "// &
       "tracer_advect_init must be called before advect_tracer.")
  if (.not. associated(Reg)) call MOM_error(FATAL, "MOM_tracer_advect: "// &
       "register_tracer must be called before advect_tracer.")
  if (Reg%ntr==0) return
  call cpu_clock_begin(id_clock_advect)

  if (.not. c_associated(CS%plan)) call make_plan(G, GV, CS)
  h = registry_hash(Reg)
  if (Reg%ntr /= CS%ntr_bound .or. h /= CS%reg_hash) then
    call check_rc(mom6adv_set_ntr(CS%plan, int(Reg%ntr, c_int)), CS%plan, "set_ntr")
    do m=1,Reg%ntr
      call check_rc(mom6adv_bind_tracer(CS%plan, int(m-1, c_int), c_loc(Reg%Tr(m)%t), &
                      real(Reg%Tr(m)%conc_underflow, c_double), ptr3(Reg%Tr(m)%ad_x), ptr3(Reg%Tr(m)%ad_y), &
                      ptr2(Reg%Tr(m)%ad2d_x), ptr2(Reg%Tr(m)%ad2d_y), ptr3(Reg%Tr(m)%advection_xy)), &
                    CS%plan, "bind_tracer")
    enddo
    CS%ntr_bound = Reg%ntr ; CS%reg_hash = h
  endif

  ! the OBC argument (:64): segment tables and tracer reservoirs go to the device on every call (tres evolves)
  call pass_obc(OBC, CS)

  ! the reference re-reads G%first_direction on every call (:126; MOM.F90:1372 toggles it under
  ! ALTERNATE_FIRST_DIRECTION), so the order is always passed explicitly
  xf = merge(1, 0, MOD(G%first_direction,2) == 0) ; if (present(x_first_in)) xf = merge(1, 0, x_first_in)
  mi = -1 ; if (present(max_iter_in)) mi = max_iter_in
  uv = -1 ; if (present(update_vol_prev)) uv = merge(1, 0, update_vol_prev)
  p_vol = c_null_ptr ; if (present(vol_prev)) p_vol = c_loc(vol_prev)
  p_uo = c_null_ptr ; if (present(uhr_out)) p_uo = c_loc(uhr_out)
  p_vo = c_null_ptr ; if (present(vhr_out)) p_vo = c_loc(vhr_out)
  call check_rc(mom6adv_advect(CS%plan, c_loc(h_end), c_loc(uhtr), c_loc(vhtr), real(dt, c_double), xf, p_vol, &
                               mi, uv, p_uo, p_vo, nit), CS%plan, "advect_tracer")

  call cpu_clock_end(id_clock_advect)
end subroutine advect_tracer

!> Marshal ocean_OBC_type into the POD tables of include/mom6adv.h and hand them to the plan.
!! Only the members advect_x / advect_y read are passed (reference :442-478,580-628,825-862,965-1014).
subroutine pass_obc(OBC, CS)
  type(ocean_OBC_type),   pointer :: OBC !< open boundary condition data (may be unassociated)
  type(tracer_advect_CS), pointer :: CS  !< control structure for this module
  type(mom6adv_obc), target :: o
  type(mom6adv_obc_segment), allocatable, target :: segs(:)
  type(mom6adv_obc_tracer),  allocatable, target :: trs(:)
  type(OBC_segment_type), pointer :: segment
  integer :: n, m, nt, off

  segment => NULL()

  if (.not. associated(OBC)) then
    if (CS%obc_was_set) call check_rc(mom6adv_set_obc(CS%plan, c_null_ptr), CS%plan, "set_obc")
    CS%obc_was_set = .false. ; return
  endif
  nt = 0
  do n=1,OBC%number_of_segments
    if (associated(OBC%segment(n)%tr_Reg)) nt = nt + OBC%segment(n)%tr_Reg%ntseg
  enddo
  allocate(segs(max(OBC%number_of_segments,1)), trs(max(nt,1)))
  off = 0
  do n=1,OBC%number_of_segments
    segment => OBC%segment(n)
    segs(n)%is_E_or_W = merge(1, 0, segment%is_E_or_W) ; segs(n)%is_N_or_S = merge(1, 0, segment%is_N_or_S)
    segs(n)%direction = segment%direction ; segs(n)%specified = merge(1, 0, segment%specified)
    segs(n)%isd = segment%HI%isd ; segs(n)%ied = segment%HI%ied ; segs(n)%jsd = segment%HI%jsd ; segs(n)%jed = segment%HI%jed
    segs(n)%IsdB = segment%HI%IsdB ; segs(n)%IedB = segment%HI%IedB
    segs(n)%JsdB = segment%HI%JsdB ; segs(n)%JedB = segment%HI%JedB
    segs(n)%ntseg = -1 ; segs(n)%Tr = c_null_ptr
    if (associated(segment%tr_Reg)) then
      segs(n)%ntseg = segment%tr_Reg%ntseg
      do m=1,segment%tr_Reg%ntseg
        trs(off+m)%ntr_index = segment%tr_Reg%Tr(m)%ntr_index
        trs(off+m)%OBC_inflow_conc = segment%tr_Reg%Tr(m)%OBC_inflow_conc
        trs(off+m)%tres = c_null_ptr
        if (allocated(segment%tr_Reg%Tr(m)%tres)) trs(off+m)%tres = c_loc(segment%tr_Reg%Tr(m)%tres)
      enddo
      if (segment%tr_Reg%ntseg > 0) segs(n)%Tr = c_loc(trs(off+1))
      off = off + segment%tr_Reg%ntseg
    endif
  enddo
  o%OBC_pe = merge(1, 0, OBC%OBC_pe) ; o%number_of_segments = OBC%number_of_segments
  o%open_u_BCs_exist_globally = merge(1, 0, OBC%open_u_BCs_exist_globally)
  o%open_v_BCs_exist_globally = merge(1, 0, OBC%open_v_BCs_exist_globally)
  o%specified_u_BCs_exist_globally = merge(1, 0, OBC%specified_u_BCs_exist_globally)
  o%specified_v_BCs_exist_globally = merge(1, 0, OBC%specified_v_BCs_exist_globally)
  o%segment = c_loc(segs(1))
  call check_rc(mom6adv_set_obc(CS%plan, c_loc(o)), CS%plan, "set_obc")
  CS%obc_was_set = .true.
  deallocate(segs, trs)
end subroutine pass_obc

!> Initialise the module: same parameters, defaults and messages as the reference (:1090-1150).
subroutine tracer_advect_init(Time, G, US, param_file, diag, CS)
  type(time_type), target, intent(in)    :: Time        !< current model time
  type(ocean_grid_type),   intent(in)    :: G           !< ocean grid structure
  type(unit_scale_type),   intent(in)    :: US          !< dimensional unit scaling type
  type(param_file_type),   intent(in)    :: param_file  !< open file to parse for model parameters
  type(diag_ctrl), target, intent(inout) :: diag        !< regulates diagnostic output
  type(tracer_advect_CS),  pointer       :: CS          !< module control structure
  ! This include declares and sets the variable "version".
# include "version_variable.h"
  character(len=40)  :: mdl = "MOM_tracer_advect" ! This module's name.
  character(len=256) :: mesg                      ! the advection scheme's name

  if (associated(CS)) then
    call MOM_error(WARNING, "tracer_advect_init called with associated control structure.")
    return
  endif
  allocate(CS)
  CS%diag => diag

  call log_version(param_file, mdl, version, "")
  call get_param(param_file, mdl, "DT", CS%dt, fail_if_missing=.true., &
                 desc="The (baroclinic) dynamics time step.", units="s", scale=US%s_to_T)
  call get_param(param_file, mdl, "DEBUG", CS%debug, default=.false.)
  ! grid connectivity for the device-side group pass (logged by MOM_domains_init already)
  call get_param(param_file, mdl, "REENTRANT_X", CS%reentrant_x, default=.true., do_not_log=.true.)
  call get_param(param_file, mdl, "REENTRANT_Y", CS%reentrant_y, default=.false., do_not_log=.true.)
  call get_param(param_file, mdl, "TRIPOLAR_N", CS%tripolar_N, default=.false., do_not_log=.true.)
  call get_param(param_file, mdl, "TRACER_ADVECTION_SCHEME", mesg, &
                 desc="The horizontal transport scheme for tracers:\n"//&
                      "  PLM    - Piecewise Linear Method\n"//&
                      "  PPM:H3 - Piecewise Parabolic Method (Huyhn 3rd order)\n"//&
                      "  PPM    - Piecewise Parabolic Method (Colella-Woodward)", default='PLM')
  select case (trim(mesg))
    case ("PLM")    ; CS%usePPM = .false.
    case ("PPM:H3") ; CS%usePPM = .true.  ; CS%useHuynh = .true.
    case ("PPM")    ; CS%usePPM = .true.  ; CS%useHuynh = .false.
    case default
      call MOM_error(FATAL, "MOM_tracer_advect, tracer_advect_init: "//&
                     "Unknown TRACER_ADVECTION_SCHEME = "//trim(mesg))
  end select
  if (.not.CS%usePPM) CS%useHuynh = .false.
  if (CS%useHuynh) then
    call get_param(param_file, mdl, "USE_HUYNH_STENCIL_BUG", CS%useHuynhStencilBug, &
                   desc="If true, use a stencil width of 2 in PPM:H3 tracer advection. This is "//&
                        "incorrect and will produce regressions in certain configurations, but may "//&
                        "be required to reproduce results in legacy simulations.", default=.false.)
  endif

  id_clock_advect = cpu_clock_id('(Ocean advect tracer)', grain=CLOCK_MODULE)
  id_clock_pass = cpu_clock_id('(Ocean tracer halo updates)', grain=CLOCK_ROUTINE)
  id_clock_sync = cpu_clock_id('(Ocean tracer global synch)', grain=CLOCK_ROUTINE)
end subroutine tracer_advect_init

!> Release the GPU plan and the control structure (:1153-1158).
!> The loop "do itt=1,num_itts ... enddo" of tracer_hordiff (MOM_tracer_hor_diff.F90:539-607) on the tracers the last
!! advect_tracer call left in HBM: call this from tracer_hordiff in place of that loop, with the routine's own khdt_x,
!! khdt_y and num_itts.  It lives in this module because the plan handle is private to it.  any_df: one of
!! df_x/df_y/df2d_x/df2d_y is associated on some tracer (not produced on the device: the call is then refused).
subroutine tracer_hordiff_iterate(CS, G, GV, khdt_x, khdt_y, num_itts, Diffuse_ML_interior, ML_KhTr_scale, any_df)
  type(tracer_advect_CS),  pointer    :: CS     !< control structure of MOM_tracer_advect (owns the GPU plan)
  type(ocean_grid_type),   intent(in) :: G      !< ocean grid structure
  type(verticalGrid_type), intent(in) :: GV     !< ocean vertical grid structure
  real, dimension(SZIB_(G),SZJ_(G)), target, intent(in) :: khdt_x !< Khtr*dt*dy/dx at u points [L2 ~> m2]
  real, dimension(SZI_(G),SZJB_(G)), target, intent(in) :: khdt_y !< Khtr*dt*dx/dy at v points [L2 ~> m2]
  integer,                 intent(in) :: num_itts            !< iterations (:368-387, after max_across_PEs)
  logical,                 intent(in) :: Diffuse_ML_interior !< CS%Diffuse_ML_interior of tracer_hor_diff_CS
  real,                    intent(in) :: ML_KhTr_scale       !< CS%ML_KhTr_scale of tracer_hor_diff_CS
  logical,                 intent(in) :: any_df              !< a diffusive-flux diagnostic is wanted
  integer(c_int) :: rc
  if (.not.associated(CS)) call MOM_error(FATAL, "tracer_hordiff_iterate: tracer_advect_init has not been called")
  if (.not.c_associated(CS%plan)) call MOM_error(FATAL, "tracer_hordiff_iterate: no advect_tracer call has made the plan yet")
  rc = mom6adv_hordiff(CS%plan, c_null_ptr, c_loc(khdt_x), c_loc(khdt_y), int(num_itts, c_int), &
                       real(GV%H_subroundoff, c_double), merge(1_c_int, 0_c_int, Diffuse_ML_interior), &
                       int(GV%nkml, c_int), int(GV%nk_rho_varies, c_int), real(ML_KhTr_scale, c_double), &
                       merge(1_c_int, 0_c_int, any_df))
  call check_rc(rc, CS%plan, "tracer_hordiff_iterate")
end subroutine tracer_hordiff_iterate

!> The part of offline_advection_ale between "Copy over the horizontal mass fluxes" and its end
!! (MOM_offline_main.F90:298-415) for runs that keep ALE out of the loop (off_ale_mod > num_off_iter): call this from
!! offline_advection_ale in place of those lines.  h_pre, uhtr and vhtr go up once and come back once; the tracers
!! stay in HBM (they were put there by the last advect_tracer call or by the registry's first binding).
subroutine offline_advection_iterate(CS, G, GV, US, h_pre, uhtr, vhtr, dt_offline, num_off_iter, off_ale_mod, &
                                     x_before_y, min_residual, converged)
  type(tracer_advect_CS),  pointer    :: CS     !< control structure of MOM_tracer_advect (owns the GPU plan)
  type(ocean_grid_type),   intent(in) :: G      !< ocean grid structure
  type(verticalGrid_type), intent(in) :: GV     !< ocean vertical grid structure
  type(unit_scale_type),   intent(in) :: US     !< a dimensional unit scaling type
  real, dimension(SZI_(G),SZJ_(G),SZK_(GV)),  target, intent(inout) :: h_pre !< layer thicknesses [H ~> m or kg m-2]
  real, dimension(SZIB_(G),SZJ_(G),SZK_(GV)), target, intent(inout) :: uhtr  !< zonal mass transport [H L2 ~> m3 or kg]
  real, dimension(SZI_(G),SZJB_(G),SZK_(GV)), target, intent(inout) :: vhtr  !< meridional mass transport [H L2 ~> m3 or kg]
  real,    intent(in)  :: dt_offline   !< CS%dt_offline of offline_transport_CS [T ~> s]
  integer, intent(in)  :: num_off_iter !< CS%num_off_iter
  integer, intent(in)  :: off_ale_mod  !< CS%off_ale_mod
  logical, intent(in)  :: x_before_y   !< CS%x_before_y
  real,    intent(in)  :: min_residual !< CS%min_residual [H L2 ~> m3 or kg]
  logical, intent(out) :: converged    !< true if no mass transport remains
  integer(c_int) :: rc, conv, nit
  integer(c_int), parameter :: F_H_END = 0, F_UHTR = 1, F_VHTR = 2, F_UHR = 4, F_VHR = 5   ! mom6adv.h field ids
  if (.not.associated(CS)) call MOM_error(FATAL, "offline_advection_iterate: tracer_advect_init has not been called")
  if (.not.c_associated(CS%plan)) call MOM_error(FATAL, "offline_advection_iterate: the GPU plan does not exist yet")
  call check_rc(mom6adv_upload_field(CS%plan, F_H_END, c_loc(h_pre)), CS%plan, "offline_advection_iterate")
  call check_rc(mom6adv_upload_field(CS%plan, F_UHTR, c_loc(uhtr)), CS%plan, "offline_advection_iterate")
  call check_rc(mom6adv_upload_field(CS%plan, F_VHTR, c_loc(vhtr)), CS%plan, "offline_advection_iterate")
  rc = mom6adv_offline_advect(CS%plan, real(dt_offline, c_double), int(num_off_iter, c_int), int(off_ale_mod, c_int), &
                              merge(1_c_int, 0_c_int, x_before_y), real(min_residual, c_double), &
                              real(GV%H_to_kg_m2, c_double), real(US%L_to_m, c_double), real(GV%kg_m2_to_H, c_double), &
                              real(US%m_to_L, c_double), conv, nit, c_null_ptr)
  call check_rc(rc, CS%plan, "offline_advection_iterate")
  converged = (conv /= 0)
  call check_rc(mom6adv_download_field(CS%plan, F_H_END, c_loc(h_pre)), CS%plan, "offline_advection_iterate")
  call check_rc(mom6adv_download_field(CS%plan, F_UHR, c_loc(uhtr)), CS%plan, "offline_advection_iterate")
  call check_rc(mom6adv_download_field(CS%plan, F_VHR, c_loc(vhtr)), CS%plan, "offline_advection_iterate")
end subroutine offline_advection_iterate

subroutine tracer_advect_end(CS)
  type(tracer_advect_CS), pointer :: CS !< module control structure
  integer(c_int) :: rc
  if (associated(CS)) then
    if (c_associated(CS%plan)) rc = mom6adv_plan_destroy(CS%plan)
    deallocate(CS)
  endif
end subroutine tracer_advect_end

end module MOM_tracer_advect
