!> Driver of the stub harness: reads a case file written by write_case.py, calls the reference's own
!! tracer_advect_init / advect_tracer (unmodified MOM_tracer_advect.F90), writes the results.
!! Test infrastructure.  File layout: see write_case.py (little-endian stream, int32 header, float64 data).
program advect_driver
use MOM_tracer_advect,   only : advect_tracer, tracer_advect_init, tracer_advect_end, tracer_advect_CS
use MOM_grid,            only : ocean_grid_type
use MOM_domains,         only : MOM_domain_type
use MOM_verticalGrid,    only : verticalGrid_type
use MOM_unit_scaling,    only : unit_scale_type
use MOM_tracer_registry, only : tracer_registry_type
use MOM_open_boundary,   only : ocean_OBC_type
use MOM_file_parser,     only : param_file_type, set_param
use MOM_diag_mediator,   only : diag_ctrl, time_type
implicit none

integer, parameter :: MAGIC = 1297041718     ! "MOM6"
type(ocean_grid_type)   :: G
type(MOM_domain_type), target :: Dom
type(verticalGrid_type) :: GV
type(unit_scale_type)   :: US
type(param_file_type)   :: pf
type(time_type), target :: Time
type(diag_ctrl), target :: diag
type(tracer_advect_CS), pointer :: CS => NULL()
type(tracer_registry_type), pointer :: Reg => NULL()
type(ocean_OBC_type), pointer :: OBC => NULL()
integer :: hdr(28), m, d, call_no, u_in, u_out, ntr, nk, scheme, x_first, max_iter_in, upd, has_vol, ncalls, diags
real :: dt, dt_bc, cu
real, allocatable, dimension(:,:,:) :: h_end, uhtr, vhtr, vol_prev, uhr_out, vhr_out
character(len=512) :: fin, fout
character(len=32) :: buf

call get_command_argument(1, fin) ; call get_command_argument(2, fout)
open(newunit=u_in, file=trim(fin), access='stream', form='unformatted', status='old')
read(u_in) hdr
if (hdr(1) /= MAGIC) error stop "not a case file"
G%isc = hdr(2) ; G%iec = hdr(3) ; G%jsc = hdr(4) ; G%jec = hdr(5)
G%isd = hdr(6) ; G%ied = hdr(7) ; G%jsd = hdr(8) ; G%jed = hdr(9)
G%IsdB = hdr(10) ; G%IedB = hdr(11) ; G%JsdB = hdr(12) ; G%JedB = hdr(13)
nk = hdr(14) ; G%ke = nk ; GV%ke = nk ; G%first_direction = hdr(15)
ntr = hdr(16) ; scheme = hdr(17) ; x_first = hdr(19) ; max_iter_in = hdr(20) ; upd = hdr(21)
has_vol = hdr(22) ; ncalls = hdr(23) ; diags = hdr(27)
Dom%cyclic_x = (hdr(24) /= 0) ; Dom%cyclic_y = (hdr(25) /= 0) ; Dom%tripolar_n = (hdr(26) /= 0)
Dom%symmetric = (G%IsdB < G%isd)
Dom%isc = G%isc ; Dom%iec = G%iec ; Dom%jsc = G%jsc ; Dom%jec = G%jec
Dom%isd = G%isd ; Dom%ied = G%ied ; Dom%jsd = G%jsd ; Dom%jed = G%jed ; Dom%IsdB = G%IsdB ; Dom%JsdB = G%JsdB
G%Domain => Dom
G%IscB = G%isc ; G%IecB = G%iec ; G%JscB = G%jsc ; G%JecB = G%jec
if (Dom%symmetric) then ; G%IscB = G%isc-1 ; G%JscB = G%jsc-1 ; endif
read(u_in) dt, dt_bc, GV%Angstrom_H, GV%H_subroundoff

allocate(G%areaT(G%isd:G%ied,G%jsd:G%jed), G%IareaT(G%isd:G%ied,G%jsd:G%jed), G%mask2dT(G%isd:G%ied,G%jsd:G%jed))
allocate(G%mask2dCu(G%IsdB:G%IedB,G%jsd:G%jed), G%mask2dCv(G%isd:G%ied,G%JsdB:G%JedB))
read(u_in) G%areaT, G%IareaT, G%mask2dT, G%mask2dCu, G%mask2dCv
allocate(h_end(G%isd:G%ied,G%jsd:G%jed,nk), uhtr(G%IsdB:G%IedB,G%jsd:G%jed,nk), vhtr(G%isd:G%ied,G%JsdB:G%JedB,nk))
allocate(uhr_out(G%IsdB:G%IedB,G%jsd:G%jed,nk), vhr_out(G%isd:G%ied,G%JsdB:G%JedB,nk))
allocate(vol_prev(G%isd:G%ied,G%jsd:G%jed,nk))
read(u_in) h_end, uhtr, vhtr
if (has_vol /= 0) read(u_in) vol_prev

allocate(Reg) ; Reg%ntr = ntr
do m=1,ntr
  read(u_in) cu
  Reg%Tr(m)%conc_underflow = cu
  allocate(Reg%Tr(m)%t(G%isd:G%ied,G%jsd:G%jed,nk))
  read(u_in) Reg%Tr(m)%t
  if (diags /= 0) then
    allocate(Reg%Tr(m)%ad_x(G%IsdB:G%IedB,G%jsd:G%jed,nk), Reg%Tr(m)%ad_y(G%isd:G%ied,G%JsdB:G%JedB,nk))
    allocate(Reg%Tr(m)%ad2d_x(G%IsdB:G%IedB,G%jsd:G%jed), Reg%Tr(m)%ad2d_y(G%isd:G%ied,G%JsdB:G%JedB))
    allocate(Reg%Tr(m)%advection_xy(G%isd:G%ied,G%jsd:G%jed,nk))
  endif
enddo
close(u_in)

write(buf,'(es24.16)') dt_bc ; call set_param(pf, "DT", buf)
if (scheme == 0) call set_param(pf, "TRACER_ADVECTION_SCHEME", "PLM")
if (scheme == 1) call set_param(pf, "TRACER_ADVECTION_SCHEME", "PPM")
if (scheme == 2) call set_param(pf, "TRACER_ADVECTION_SCHEME", "PPM:H3")
if (hdr(18) /= 0) call set_param(pf, "USE_HUYNH_STENCIL_BUG", "True")
call tracer_advect_init(Time, G, US, pf, diag, CS)

open(newunit=u_out, file=trim(fout), access='stream', form='unformatted', status='replace')
do call_no=1,ncalls
  ! every combination of absent / present optionals the case files use
  if (has_vol /= 0) then
    if (x_first >= 0) then
      call advect_tracer(h_end, uhtr, vhtr, OBC, dt, G, GV, US, CS, Reg, x_first_in=(x_first/=0), vol_prev=vol_prev, &
                         max_iter_in=max_iter_in, update_vol_prev=(upd>0), uhr_out=uhr_out, vhr_out=vhr_out)
    else
      call advect_tracer(h_end, uhtr, vhtr, OBC, dt, G, GV, US, CS, Reg, vol_prev=vol_prev, &
                         max_iter_in=max_iter_in, update_vol_prev=(upd>0), uhr_out=uhr_out, vhr_out=vhr_out)
    endif
  elseif (x_first >= 0) then
    call advect_tracer(h_end, uhtr, vhtr, OBC, dt, G, GV, US, CS, Reg, x_first_in=(x_first/=0), &
                       uhr_out=uhr_out, vhr_out=vhr_out)
  else
    call advect_tracer(h_end, uhtr, vhtr, OBC, dt, G, GV, US, CS, Reg, uhr_out=uhr_out, vhr_out=vhr_out)
  endif
enddo
do m=1,ntr
  write(u_out) Reg%Tr(m)%t
enddo
write(u_out) uhr_out, vhr_out
if (has_vol /= 0) write(u_out) vol_prev
if (diags /= 0) then
  do m=1,ntr
    write(u_out) Reg%Tr(m)%ad_x, Reg%Tr(m)%ad_y, Reg%Tr(m)%ad2d_x, Reg%Tr(m)%ad2d_y, Reg%Tr(m)%advection_xy
  enddo
endif
close(u_out)
call tracer_advect_end(CS)
end program advect_driver
