! gpu_interface.f90 -- iso_c_binding shim between critic2's Fortran drivers and libc2gpu (include/c2gpu.h).
!
! Sits in src/ next to c_interface_module.f90 and follows the in-tree Fortran->C pattern
! (interface ... bind(c) blocks, src/crystalmod@proc.f90:3170-3186; implementation in C, src/doqhull.c).
! The host program stays Fortran: this module only flattens type(molwfn) / type(environ) / agrid(:) into the
! plain arrays of the C ABI and forwards the grid descriptors of the three evaluation loops it replaces:
!   rhoplot_cube        src/rhoplot@proc.f90:582-631
!   nciplot             src/nci@proc.f90:499-561 (evaluation) and :569-604 (cutoff pass)
!   promolecular_grid   src/crystalmod@proc.f90:5118-5158
! Status convention: 0 ok; -1 unsupported input (the caller keeps its OpenMP loop); >0 error -> ferror(faterr).
!
! NOT compiled in the authoring container (no Fortran compiler exists there); see INTEGRATION.md.
module gpu_interface
  use iso_c_binding, only: c_int, c_long, c_double, c_ptr, c_char, c_null_ptr, c_null_char, c_associated, c_loc
  implicit none
  private

  public :: gpu_enabled, gpu_init, gpu_end
  public :: gpu_wfn_handle, gpu_promol_handle, gpu_set_core, gpu_destroy
  public :: gpu_cube, gpu_nciplot, gpu_nci_cutoffs, gpu_promolecular_grid, gpu_fillmesh, gpu_fillmesh_mo
  public :: gpu_fragment_handle, gpu_nci_fragments, gpu_nci_cutoffs_frag
  public :: gpu_write_cube_body, gpu_write_dat

  logical, save :: gpu_enabled = .false. !< set by the GPU keyword (global@proc.f90, critic_setvariables)

  interface
     function c2gpu_init(ndev_req,ndev) bind(c,name="c2gpu_init") result(ierr)
       import :: c_int
       integer(c_int), value :: ndev_req
       integer(c_int) :: ndev
       integer(c_int) :: ierr
     end function c2gpu_init
     function c2gpu_finalize() bind(c,name="c2gpu_finalize") result(ierr)
       import :: c_int
       integer(c_int) :: ierr
     end function c2gpu_finalize
     function c2gpu_last_error(buf,len) bind(c,name="c2gpu_last_error") result(ierr)
       import :: c_int, c_char
       character(kind=c_char) :: buf(*)
       integer(c_int), value :: len
       integer(c_int) :: ierr
     end function c2gpu_last_error
     function c2gpu_wfn_create(natom,xyz,npri,icenter,itype,expo,dran,nmoocc,ldc,cmo,occ,wfntyp,nalpha,&
        nedf,icenter_edf,itype_edf,e_edf,c_edf,dran_edf,ismolecule,m_x2c,m_c2x,handle) &
        bind(c,name="c2gpu_wfn_create") result(ierr)
       import :: c_int, c_double, c_ptr
       integer(c_int), value :: natom, npri, nmoocc, ldc, wfntyp, nalpha, nedf, ismolecule
       real(c_double) :: xyz(3,*), expo(*), dran(*), cmo(ldc,*), occ(*), e_edf(*), c_edf(*), dran_edf(*)
       integer(c_int) :: icenter(*), itype(*), icenter_edf(*), itype_edf(*)
       real(c_double) :: m_x2c(3,3), m_c2x(3,3)
       type(c_ptr) :: handle
       integer(c_int) :: ierr
     end function c2gpu_wfn_create
     function c2gpu_promol_create(nenv,ncell,xyz_env,ispc,cidx,nspc,zspc,ngrid,a,b,cutrad,tab_r,tab_f,tab_fp,&
        tab_fpp,ismolecule,m_x2c,m_c2x,m_xr2c,m_c2xr,fragmask,handle) bind(c,name="c2gpu_promol_create") result(ierr)
       import :: c_int, c_double, c_ptr
       integer(c_int), value :: nenv, ncell, nspc, ismolecule
       real(c_double) :: xyz_env(3,*), a(*), b(*), cutrad(*), tab_r(*), tab_f(*), tab_fp(*), tab_fpp(*)
       integer(c_int) :: ispc(*), cidx(*), zspc(*), ngrid(*)
       real(c_double) :: m_x2c(3,3), m_c2x(3,3), m_xr2c(3,3), m_c2xr(3,3)
       type(c_ptr), value :: fragmask
       type(c_ptr) :: handle
       integer(c_int) :: ierr
     end function c2gpu_promol_create
     function c2gpu_destroy(handle) bind(c,name="c2gpu_destroy") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: handle
       integer(c_int) :: ierr
     end function c2gpu_destroy
     function c2gpu_field_set_core(handle,hcore) bind(c,name="c2gpu_field_set_core") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: handle, hcore
       integer(c_int) :: ierr
     end function c2gpu_field_set_core
     function c2gpu_eval_grid(h,x0,xd,nn,nder,periodic,outmask,out) bind(c,name="c2gpu_eval_grid") result(ierr)
       import :: c_int, c_double, c_ptr
       type(c_ptr), value :: h
       real(c_double) :: x0(3), xd(3,3)
       integer(c_int) :: nn(3)
       integer(c_int), value :: nder, periodic, outmask
       type(c_ptr) :: out(*)
       integer(c_int) :: ierr
     end function c2gpu_eval_grid
     function c2gpu_nci_grid(href,hpromol,x0,xmat,nstep,crho,cgrad,rhoat) bind(c,name="c2gpu_nci_grid") result(ierr)
       import :: c_int, c_double, c_ptr
       type(c_ptr), value :: href, hpromol
       real(c_double) :: x0(3), xmat(3,3)
       integer(c_int) :: nstep(3)
       real(c_double) :: crho(*), cgrad(*)
       type(c_ptr), value :: rhoat
       integer(c_int) :: ierr
     end function c2gpu_nci_grid
     function c2gpu_nci_cutoffs(npts,crho,cgrad,rhoat,rhocut,dimcut,rhoplot,srhorange,onlyneg,rho_void,omega,&
        sums,ndat,dat,maxdat) bind(c,name="c2gpu_nci_cutoffs") result(ierr)
       import :: c_int, c_long, c_double, c_ptr
       integer(c_long), value :: npts, maxdat
       real(c_double) :: crho(*), cgrad(*)
       type(c_ptr), value :: rhoat, dat
       real(c_double), value :: rhocut, dimcut, rhoplot, rho_void, omega
       real(c_double) :: srhorange(2), sums(3)
       integer(c_int), value :: onlyneg
       integer(c_long) :: ndat
       integer(c_int) :: ierr
     end function c2gpu_nci_cutoffs
     function c2gpu_nci_fragments(nfrag,hfrag,x0,xmat,nstep,rhofrag) bind(c,name="c2gpu_nci_fragments") result(ierr)
       import :: c_int, c_double, c_ptr
       integer(c_int), value :: nfrag
       type(c_ptr) :: hfrag(*)
       real(c_double) :: x0(3), xmat(3,3)
       integer(c_int) :: nstep(3)
       real(c_double) :: rhofrag(*)
       integer(c_int) :: ierr
     end function c2gpu_nci_fragments
     function c2gpu_nci_cutoffs_frag(npts,crho,cgrad,rhoat,nfrag,rhofrag,rhoparam,rhoparam2,rhocut,dimcut,rhoplot,&
        srhorange,onlyneg,rho_void,omega,sums,ndat,dat,maxdat) bind(c,name="c2gpu_nci_cutoffs_frag") result(ierr)
       import :: c_int, c_long, c_double, c_ptr
       integer(c_long), value :: npts, maxdat
       real(c_double) :: crho(*), cgrad(*)
       type(c_ptr), value :: rhoat, rhofrag, dat
       integer(c_int), value :: nfrag, onlyneg
       real(c_double), value :: rhoparam, rhoparam2, rhocut, dimcut, rhoplot, rho_void, omega
       real(c_double) :: srhorange(2), sums(3)
       integer(c_long) :: ndat
       integer(c_int) :: ierr
     end function c2gpu_nci_cutoffs_frag
     function c2gpu_cube_text_size(n,fmt) bind(c,name="c2gpu_cube_text_size") result(nbytes)
       import :: c_int, c_long
       integer(c_int) :: n(3)
       integer(c_int), value :: fmt
       integer(c_long) :: nbytes
     end function c2gpu_cube_text_size
     function c2gpu_cube_text(g,n,fmt,order,text) bind(c,name="c2gpu_cube_text") result(ierr)
       import :: c_int, c_double, c_char
       real(c_double) :: g(*)
       integer(c_int) :: n(3)
       integer(c_int), value :: fmt, order
       character(kind=c_char) :: text(*)
       integer(c_int) :: ierr
     end function c2gpu_cube_text
     function c2gpu_dat_text_size(npairs) bind(c,name="c2gpu_dat_text_size") result(nbytes)
       import :: c_long
       integer(c_long), value :: npairs
       integer(c_long) :: nbytes
     end function c2gpu_dat_text_size
     function c2gpu_dat_text(npairs,dat,text) bind(c,name="c2gpu_dat_text") result(ierr)
       import :: c_int, c_long, c_double, c_char
       integer(c_long), value :: npairs
       real(c_double) :: dat(2,*)
       character(kind=c_char) :: text(*)
       integer(c_int) :: ierr
     end function c2gpu_dat_text
     function c2gpu_eval_points(h,npts,pts,nder,periodic,res) bind(c,name="c2gpu_eval_points") result(ierr)
       import :: c_int, c_long, c_double, c_ptr
       type(c_ptr), value :: h
       integer(c_long), value :: npts
       real(c_double), intent(in) :: pts(3,*)
       integer(c_int), value :: nder, periodic
       real(c_double), intent(inout) :: res(16,*)
       integer(c_int) :: ierr
     end function c2gpu_eval_points
     function c2gpu_eval_points_gkin(h,npts,pts,periodic,gkin) bind(c,name="c2gpu_eval_points_gkin") result(ierr)
       import :: c_int, c_long, c_double, c_ptr
       type(c_ptr), value :: h
       integer(c_long), value :: npts
       real(c_double), intent(in) :: pts(3,*)
       integer(c_int), value :: periodic
       real(c_double), intent(inout) :: gkin(*)
       integer(c_int) :: ierr
     end function c2gpu_eval_points_gkin
     function c2gpu_eval_points_mo(hmo,npts,pts,periodic,phi) bind(c,name="c2gpu_eval_points_mo") result(ierr)
       import :: c_int, c_long, c_double, c_ptr
       type(c_ptr), value :: hmo
       integer(c_long), value :: npts
       real(c_double), intent(in) :: pts(3,*)
       integer(c_int), value :: periodic
       real(c_double), intent(inout) :: phi(*)
       integer(c_int) :: ierr
     end function c2gpu_eval_points_mo
     function c2gpu_promol_grid(h,n,out) bind(c,name="c2gpu_promol_grid") result(ierr)
       import :: c_int, c_double, c_ptr
       type(c_ptr), value :: h
       integer(c_int) :: n(3)
       real(c_double) :: out(*)
       integer(c_int) :: ierr
     end function c2gpu_promol_grid
  end interface

contains

  !> Turn a non-zero status into the reference's error convention (tools_io ferror).  -1 is not an error.
  subroutine gpu_check(routine,ierr)
    use tools_io, only: ferror, faterr
    character*(*), intent(in) :: routine
    integer(c_int), intent(in) :: ierr
    character(kind=c_char) :: buf(512)
    character(len=512) :: msg
    integer :: i, idum
    if (ierr <= 0) return
    idum = c2gpu_last_error(buf,512_c_int)
    msg = ""
    do i = 1, 512
       if (buf(i) == c_null_char) exit
       msg(i:i) = buf(i)
    end do
    call ferror(routine,"GPU module: " // trim(msg),faterr)
  end subroutine gpu_check

  subroutine gpu_init(ndev)
    integer, intent(out) :: ndev
    integer(c_int) :: n, ierr
    ierr = c2gpu_init(0_c_int,n)
    call gpu_check("gpu_init",ierr)
    ndev = n
    gpu_enabled = (ierr == 0 .and. n > 0)
  end subroutine gpu_init

  subroutine gpu_end()
    integer(c_int) :: ierr
    ierr = c2gpu_finalize()
    gpu_enabled = .false.
  end subroutine gpu_end

  !> Device image of a wavefunction field (called from field_new after complete_struct; freed in field_end).
  !> Returns c_null_ptr when the input is unsupported (STO basis, crystal environment): the caller keeps grd.
  !> With imo the handle holds that molecular orbital alone (nmoocc = 1, cmo = cmo(imo,:), occ = 1): the operand of
  !> c2gpu_eval_points_mo (calculate_mo, wfn_private@proc.F90:2447-2553; imo may address a virtual orbital).
  !> With spin = 1 / 2 and/or valence = .true. the handle is one channel of rho2 (wfn_private@proc.F90:2091-2147,
  !> 2159-2200): the alpha / beta MOs of a UHF wavefunction plus half the EDF core (half of everything otherwise), and/or
  !> the density without the EDF core -- the operand behind res%fspin, gfmodspin, lapspin, gkinspin and res%fval,
  !> gfmodval, del2fval (fieldmod@proc.f90:763-786).  Only the occupations and EDF coefficients passed differ.
  function gpu_wfn_handle(w,c,imo,spin,valence) result(h)
    use wfn_private, only: molwfn, wfn_frac, wfn_uhf
    use crystalmod, only: crystal
    type(molwfn), intent(in), target :: w
    type(crystal), intent(in) :: c
    integer, intent(in), optional :: imo
    integer, intent(in), optional :: spin
    logical, intent(in), optional :: valence
    type(c_ptr) :: h
    real(c_double), allocatable :: occ(:), cedf(:)
    integer :: ispin, nedf
    real(c_double), allocatable :: row(:)
    real(c_double) :: one(1)
    real(c_double), allocatable :: xyz(:,:)
    integer(c_int) :: ierr, ldc
    integer :: i
    real(c_double) :: dumr(1)
    integer(c_int) :: dumi(1)
    h = c_null_ptr
    if (w%issto .or. .not.c%ismolecule) return
    allocate(xyz(3,w%env%ncell))
    do i = 1, w%env%ncell
       xyz(:,i) = w%env%at(i)%r
    end do
    ldc = size(w%cmo,1)
    if (present(imo)) then
       if (imo < 1 .or. imo > ldc) return
       allocate(row(w%npri))
       row = w%cmo(imo,:)
       one = 1d0
       ierr = c2gpu_wfn_create(int(w%env%ncell,c_int),xyz,int(w%npri,c_int),w%icenter,w%itype,w%e,w%dran,&
          1_c_int,1_c_int,row,one,0_c_int,1_c_int,0_c_int,dumi,dumi,dumr,dumr,dumr,1_c_int,c%m_x2c,c%m_c2x,h)
    else
       ! occupations and EDF coefficients of the requested channel (total: as they are)
       ispin = 0
       if (present(spin)) ispin = spin
       allocate(occ(w%nmoocc))
       occ = w%occ(1:w%nmoocc)
       nedf = w%nedf
       if (present(valence)) then
          if (valence) nedf = 0
       end if
       if (ispin == 1 .or. ispin == 2) then
          if (w%wfntyp == wfn_uhf) then
             if (ispin == 1) then
                occ(w%nalpha+1:) = 0d0
             else
                occ(1:w%nalpha) = 0d0
             end if
          else
             occ = 0.5d0 * occ
          end if
       end if
       if (nedf > 0) then
          allocate(cedf(nedf))
          cedf = w%c_edf(1:nedf)
          if (ispin == 1 .or. ispin == 2) cedf = 0.5d0 * cedf
          ierr = c2gpu_wfn_create(int(w%env%ncell,c_int),xyz,int(w%npri,c_int),w%icenter,w%itype,w%e,w%dran,&
             int(w%nmoocc,c_int),ldc,w%cmo,occ,int(w%wfntyp,c_int),int(w%nalpha,c_int),int(nedf,c_int),&
             w%icenter_edf,w%itype_edf,w%e_edf,cedf,w%dran_edf,1_c_int,c%m_x2c,c%m_c2x,h)
       else
          ierr = c2gpu_wfn_create(int(w%env%ncell,c_int),xyz,int(w%npri,c_int),w%icenter,w%itype,w%e,w%dran,&
             int(w%nmoocc,c_int),ldc,w%cmo,occ,int(w%wfntyp,c_int),int(w%nalpha,c_int),0_c_int,&
             dumi,dumi,dumr,dumr,dumr,1_c_int,c%m_x2c,c%m_c2x,h)
       end if
    end if
    if (ierr == -1) h = c_null_ptr
    call gpu_check("gpu_wfn_handle",ierr)
  end function gpu_wfn_handle

  !> Device image of the promolecular field f(0): environment atoms + agrid(:) radial tables.
  !> With zpsp (f%zpsp of a field loaded with CORE/ZPSP) the tables are the core selection cgrid(iz,zpsp(iz)) that
  !> environ%promolecular uses when iscore (environmod@proc.f90:1252-1253,1295-1299): the handle is then the core
  !> field to attach with gpu_set_core.
  function gpu_promol_handle(c,fr,zpsp) result(h)
    use crystalmod, only: crystal
    use fragmentmod, only: fragment
    use grid1mod, only: agrid, cgrid, grid1
    use global, only: cutrad
    use param, only: maxzat
    type(crystal), intent(in) :: c
    type(fragment), intent(in), optional :: fr
    integer, intent(in), optional :: zpsp(:)
    type(grid1), pointer :: g
    type(c_ptr) :: h
    real(c_double), allocatable :: xyz(:,:), a(:), b(:), cut(:), tr(:), tf(:), tfp(:), tfpp(:)
    integer(c_int), allocatable :: ispc(:), cidx(:), zspc(:), ngrid(:)
    character(kind=c_char), allocatable, target :: mask(:)
    type(c_ptr) :: pmask
    integer :: i, is, iz, ntot, n0
    integer(c_int) :: ierr
    h = c_null_ptr
    pmask = c_null_ptr
    allocate(xyz(3,c%env%n),ispc(c%env%n),cidx(c%env%n))
    do i = 1, c%env%n
       xyz(:,i) = c%env%at(i)%r
       ispc(i) = c%env%at(i)%is
       cidx(i) = c%env%at(i)%cidx
    end do
    allocate(zspc(c%nspc),ngrid(c%nspc),a(c%nspc),b(c%nspc),cut(c%nspc))
    ntot = 0
    do is = 1, c%nspc
       iz = c%spc(is)%z
       zspc(is) = iz
       ngrid(is) = 0; a(is) = 0d0; b(is) = 1d0; cut(is) = 0d0
       if (iz > 0 .and. iz <= maxzat) then
          cut(is) = cutrad(iz)
          g => promol_table(iz)
          if (associated(g)) then
             ngrid(is) = g%ngrid; a(is) = g%a; b(is) = g%b
          end if
       end if
       ntot = ntot + ngrid(is)
    end do
    allocate(tr(max(ntot,1)),tf(max(ntot,1)),tfp(max(ntot,1)),tfpp(max(ntot,1)))
    n0 = 0
    do is = 1, c%nspc
       if (ngrid(is) == 0) cycle
       iz = zspc(is)
       g => promol_table(iz)
       tr(n0+1:n0+ngrid(is)) = g%r(1:ngrid(is))
       tf(n0+1:n0+ngrid(is)) = g%f(1:ngrid(is))
       tfp(n0+1:n0+ngrid(is)) = g%fp(1:ngrid(is))
       tfpp(n0+1:n0+ngrid(is)) = g%fpp(1:ngrid(is))
       n0 = n0 + ngrid(is)
    end do
    if (present(fr)) then
       ! type_promol_frag: one byte per cell atom, 1 = atom belongs to the fragment (environmod@proc.f90:1287-1290)
       allocate(mask(c%ncel))
       mask = achar(0)
       do i = 1, fr%nat
          mask(fr%at(i)%cidx) = achar(1)
       end do
       pmask = c_loc(mask(1))
    end if
    ierr = c2gpu_promol_create(int(c%env%n,c_int),int(c%ncel,c_int),xyz,ispc,cidx,int(c%nspc,c_int),zspc,ngrid,&
       a,b,cut,tr,tf,tfp,tfpp,merge(1_c_int,0_c_int,c%ismolecule),c%m_x2c,c%m_c2x,c%m_xr2c,c%m_c2xr,pmask,h)
    if (ierr == -1) h = c_null_ptr
    call gpu_check("gpu_promol_handle",ierr)
  contains
    !> agrid(iz), or cgrid(iz,zpsp(iz)) for the core field; null() if not initialised or without electrons
    !> (interp returns 0 when z - qat <= 0, grid1mod@proc.f90:100)
    function promol_table(iz) result(g)
      integer, intent(in) :: iz
      type(grid1), pointer :: g
      g => null()
      if (present(zpsp)) then
         if (.not.allocated(cgrid)) return
         if (zpsp(iz) <= 0 .or. zpsp(iz) >= iz) return
         if (cgrid(iz,zpsp(iz))%isinit) g => cgrid(iz,zpsp(iz))
      else
         if (agrid(iz)%isinit) g => agrid(iz)
      end if
    end function promol_table
  end function gpu_promol_handle

  !> f%usecore (fieldmod@proc.f90:827-832, :922-925): attach the core field (gpu_promol_handle(c,zpsp=f%zpsp)) to the
  !> handle of field f; every GPU driver then adds the core density and derivatives before the tail of grd.
  !> Call from field_new/field_set_options after the handle exists, when f%usecore .and. any(f%zpsp > 0).
  subroutine gpu_set_core(h,hcore)
    type(c_ptr), intent(in) :: h, hcore
    call gpu_check("gpu_set_core",c2gpu_field_set_core(h,hcore))
  end subroutine gpu_set_core

  subroutine gpu_destroy(h)
    type(c_ptr), intent(inout) :: h
    integer(c_int) :: ierr
    if (c_associated(h)) ierr = c2gpu_destroy(h)
    h = c_null_ptr
  end subroutine gpu_destroy

  !> Replacement of the rhoplot_cube evaluation loop (rhoplot@proc.f90:587-631) for ONE property nti.
  !> ok = .false. means "unsupported": the caller runs its own loop.
  subroutine gpu_cube(h,x0,xd,nn,periodic,nti,lf,ok)
    type(c_ptr), intent(in) :: h
    real(c_double), intent(in) :: x0(3), xd(3,3)
    integer, intent(in) :: nn(3), nti   !< nti: 0 f, 1 gx, 2 gy, 3 gz, 4 gmod, 5 hxx ... 10 hzz, 11 lap (:488-523)
    logical, intent(in) :: periodic
    real(c_double), intent(inout), target :: lf(:,:,:)
    logical, intent(out) :: ok
    type(c_ptr) :: out(1)
    integer(c_int) :: ierr, n(3)
    ok = .false.
    if (.not.gpu_enabled .or. .not.c_associated(h)) return
    n = int(nn,c_int)
    out(1) = c_loc(lf(1,1,1))
    ierr = c2gpu_eval_grid(h,x0,xd,n,2_c_int,merge(1_c_int,0_c_int,periodic),int(ishft(1,nti),c_int),out)
    call gpu_check("rhoplot_cube",ierr)
    ok = (ierr == 0)
  end subroutine gpu_cube

  !> Replacement of the nciplot evaluation loop (nci@proc.f90:499-561); crho/cgrad(0:n3-1,0:n2-1,0:n1-1).
  subroutine gpu_nciplot(href,hpromol,x0,xmat,nstep,crho,cgrad,rhoat,ok)
    type(c_ptr), intent(in) :: href, hpromol
    real(c_double), intent(in) :: x0(3), xmat(3,3)
    integer, intent(in) :: nstep(3)
    real(c_double), intent(inout) :: crho(0:,0:,0:), cgrad(0:,0:,0:)
    real(c_double), intent(inout), target, optional :: rhoat(0:,0:,0:)
    logical, intent(out) :: ok
    integer(c_int) :: ierr, n(3)
    type(c_ptr) :: pat
    ok = .false.
    if (.not.gpu_enabled .or. .not.c_associated(href)) return
    n = int(nstep,c_int)
    pat = c_null_ptr
    if (present(rhoat) .and. c_associated(hpromol)) pat = c_loc(rhoat(0,0,0))
    ierr = c2gpu_nci_grid(href,hpromol,x0,xmat,n,crho,cgrad,pat)
    call gpu_check("nciplot",ierr)
    ok = (ierr == 0)
  end subroutine gpu_nciplot

  !> Replacement of the nciplot cutoff pass (nci@proc.f90:569-604, nfrag = 0): void sums, .dat pairs, cgrad masking.
  subroutine gpu_nci_cutoffs(crho,cgrad,rhoat,rhocut,dimcut,rhoplot,srhorange,onlyneg,rho_void,omega,&
     sumq,sumqp,sumv,ndat,dat,ok)
    real(c_double), intent(inout) :: crho(0:,0:,0:), cgrad(0:,0:,0:)
    real(c_double), intent(in), target, optional :: rhoat(0:,0:,0:)
    real(c_double), intent(in) :: rhocut, dimcut, rhoplot, srhorange(2), rho_void, omega
    logical, intent(in) :: onlyneg
    real(c_double), intent(out) :: sumq, sumqp, sumv
    integer(c_long), intent(out) :: ndat
    real(c_double), intent(inout), target :: dat(:,:)   !< (2,maxdat): rho, s pairs of the .dat file (:593-595)
    logical, intent(out) :: ok
    real(c_double) :: sums(3)
    integer(c_int) :: ierr
    type(c_ptr) :: pat
    ok = .false.
    if (.not.gpu_enabled) return
    pat = c_null_ptr
    if (present(rhoat)) pat = c_loc(rhoat(0,0,0))
    ierr = c2gpu_nci_cutoffs(int(size(crho),c_long),crho,cgrad,pat,rhocut,dimcut,rhoplot,srhorange,&
       merge(1_c_int,0_c_int,onlyneg),rho_void,omega,sums,ndat,c_loc(dat(1,1)),int(size(dat,2),c_long))
    call gpu_check("nciplot",ierr)
    sumq = sums(1); sumqp = sums(2); sumv = sums(3)
    ok = (ierr == 0)
  end subroutine gpu_nci_cutoffs

  !> Device image of one NCIPLOT fragment (nci@proc.f90:534-550): the fragment's own atoms, cutoff agrid(iz)%rmax only
  !> (cutrad = 1e30), molecule frame with identity matrices so that the point is used as is (:538).
  function gpu_fragment_handle(fr) result(h)
    use fragmentmod, only: fragment
    use grid1mod, only: agrid
    use param, only: maxzat
    type(fragment), intent(in) :: fr
    type(c_ptr) :: h
    real(c_double), allocatable :: xyz(:,:), a(:), b(:), cut(:), tr(:), tf(:)
    integer(c_int), allocatable :: ispc(:), cidx(:), zspc(:), ngrid(:)
    real(c_double) :: eye(3,3)
    integer :: i, is, iz, ntot, n0
    integer(c_int) :: ierr
    h = c_null_ptr
    if (fr%nat == 0) return
    eye = 0d0
    do i = 1, 3
       eye(i,i) = 1d0
    end do
    allocate(xyz(3,fr%nat),ispc(fr%nat),cidx(fr%nat))
    do i = 1, fr%nat
       xyz(:,i) = fr%at(i)%r
       ispc(i) = fr%at(i)%is
       cidx(i) = i
    end do
    allocate(zspc(fr%nspc),ngrid(fr%nspc),a(fr%nspc),b(fr%nspc),cut(fr%nspc))
    cut = 1d30
    ntot = 0
    do is = 1, fr%nspc
       iz = fr%spc(is)%z
       zspc(is) = iz
       ngrid(is) = 0; a(is) = 0d0; b(is) = 1d0
       if (iz > 0 .and. iz <= maxzat) then
          if (agrid(iz)%isinit) then
             ngrid(is) = agrid(iz)%ngrid; a(is) = agrid(iz)%a; b(is) = agrid(iz)%b
          end if
       end if
       ntot = ntot + ngrid(is)
    end do
    allocate(tr(max(ntot,1)),tf(max(ntot,1)))
    n0 = 0
    do is = 1, fr%nspc
       if (ngrid(is) == 0) cycle
       iz = zspc(is)
       tr(n0+1:n0+ngrid(is)) = agrid(iz)%r(1:ngrid(is))
       tf(n0+1:n0+ngrid(is)) = agrid(iz)%f(1:ngrid(is))
       n0 = n0 + ngrid(is)
    end do
    ! only the value table is read (nder = 0): f stands in for fp and fpp
    ierr = c2gpu_promol_create(int(fr%nat,c_int),int(fr%nat,c_int),xyz,ispc,cidx,int(fr%nspc,c_int),zspc,ngrid,&
       a,b,cut,tr,tf,tf,tf,1_c_int,eye,eye,eye,eye,c_null_ptr,h)
    if (ierr == -1) h = c_null_ptr
    call gpu_check("gpu_fragment_handle",ierr)
  end function gpu_fragment_handle

  !> Replacement of the fragment part of the nciplot loop (:534-550, :556): rhofrag(0:n3-1,0:n2-1,0:n1-1,nfrag).
  subroutine gpu_nci_fragments(hfrag,x0,xmat,nstep,rhofrag,ok)
    type(c_ptr), intent(in) :: hfrag(:)
    real(c_double), intent(in) :: x0(3), xmat(3,3)
    integer, intent(in) :: nstep(3)
    real(c_double), intent(inout) :: rhofrag(0:,0:,0:,1:)
    logical, intent(out) :: ok
    integer(c_int) :: ierr, n(3)
    type(c_ptr) :: hh(size(hfrag))
    ok = .false.
    if (.not.gpu_enabled) return
    n = int(nstep,c_int)
    hh = hfrag
    ierr = c2gpu_nci_fragments(int(size(hfrag),c_int),hh,x0,xmat,n,rhofrag)
    call gpu_check("nciplot",ierr)
    ok = (ierr == 0)
  end subroutine gpu_nci_fragments

  !> Cutoff pass with fragments (nci@proc.f90:569-604 incl. the inter test :579-582).
  subroutine gpu_nci_cutoffs_frag(crho,cgrad,rhoat,rhofrag,rhoparam,rhoparam2,rhocut,dimcut,rhoplot,srhorange,&
     onlyneg,rho_void,omega,sumq,sumqp,sumv,ndat,dat,ok)
    real(c_double), intent(inout) :: crho(0:,0:,0:), cgrad(0:,0:,0:)
    real(c_double), intent(in), target :: rhoat(0:,0:,0:), rhofrag(0:,0:,0:,1:)
    real(c_double), intent(in) :: rhoparam, rhoparam2, rhocut, dimcut, rhoplot, srhorange(2), rho_void, omega
    logical, intent(in) :: onlyneg
    real(c_double), intent(out) :: sumq, sumqp, sumv
    integer(c_long), intent(out) :: ndat
    real(c_double), intent(inout), target :: dat(:,:)
    logical, intent(out) :: ok
    real(c_double) :: sums(3)
    integer(c_int) :: ierr
    ok = .false.
    if (.not.gpu_enabled) return
    ierr = c2gpu_nci_cutoffs_frag(int(size(crho),c_long),crho,cgrad,c_loc(rhoat(0,0,0)),int(size(rhofrag,4),c_int),&
       c_loc(rhofrag(0,0,0,1)),rhoparam,rhoparam2,rhocut,dimcut,rhoplot,srhorange,merge(1_c_int,0_c_int,onlyneg),&
       rho_void,omega,sums,ndat,c_loc(dat(1,1)),int(size(dat,2),c_long))
    call gpu_check("nciplot",ierr)
    sumq = sums(1); sumqp = sums(2); sumv = sums(3)
    ok = (ierr == 0)
  end subroutine gpu_nci_cutoffs_frag

  !> Body of a cube file written through a stream-access unit: replaces the formatted loops of writegrid_cube
  !> (crystalmod@proc.f90:5026-5036; fmt 0 = precisecube, 1 = default, order 0: g(ix,iy,iz)) and write_cube_body
  !> (nci@proc.f90:867-883; fmt 2, order 1: c(k,j,i)).  The header is written by the caller as today; the unit is then
  !> re-opened with access='stream', position='append' and the text block written in one statement.
  subroutine gpu_write_cube_body(file,g,n,fmt,order,ok)
    character*(*), intent(in) :: file
    real(c_double), intent(in) :: g(*)
    integer, intent(in) :: n(3), fmt, order
    logical, intent(out) :: ok
    character(kind=c_char), allocatable :: text(:)
    integer(c_int) :: ierr, nn(3)
    integer(c_long) :: nbytes
    integer :: lu
    ok = .false.
    if (.not.gpu_enabled) return
    nn = int(n,c_int)
    nbytes = c2gpu_cube_text_size(nn,int(fmt,c_int))
    if (nbytes <= 0) return
    allocate(text(nbytes))
    ierr = c2gpu_cube_text(g,nn,int(fmt,c_int),int(order,c_int),text)
    call gpu_check("write_cube_body",ierr)
    if (ierr /= 0) return
    open(newunit=lu,file=file,access='stream',form='unformatted',position='append',status='old')
    write(lu) text
    close(lu)
    ok = .true.
  end subroutine gpu_write_cube_body

  !> The (rho,s) lines of the NCIPLOT .dat file, '(1p,E15.7,E15.7)' (nci@proc.f90:593-595), appended to `file` through a
  !> stream-access unit like gpu_write_cube_body: dat(2,ndat) are the pairs gpu_nci_cutoffs / gpu_nci_cutoffs_frag return.
  subroutine gpu_write_dat(file,dat,ndat,ok)
    character*(*), intent(in) :: file
    real(c_double), intent(in) :: dat(2,*)
    integer, intent(in) :: ndat
    logical, intent(out) :: ok
    character(kind=c_char), allocatable :: text(:)
    integer(c_int) :: ierr
    integer(c_long) :: nbytes
    integer :: lu
    ok = .false.
    if (.not.gpu_enabled) return
    if (ndat <= 0) then
       ok = .true.
       return
    end if
    nbytes = c2gpu_dat_text_size(int(ndat,c_long))
    if (nbytes <= 0) return
    allocate(text(nbytes))
    ierr = c2gpu_dat_text(int(ndat,c_long),dat,text)
    call gpu_check("write_dat",ierr)
    if (ierr /= 0) return
    open(newunit=lu,file=file,access='stream',form='unformatted',position='append',status='old')
    write(lu) text
    close(lu)
    ok = .true.
  end subroutine gpu_write_dat

  !> Replacement of the promolecular_grid loop (crystalmod@proc.f90:5143-5158): f(n1,n2,n3), nder = 0.
  subroutine gpu_promolecular_grid(h,n,f,ok)
    type(c_ptr), intent(in) :: h
    integer, intent(in) :: n(3)
    real(c_double), intent(inout) :: f(:,:,:)
    logical, intent(out) :: ok
    integer(c_int) :: ierr, nn(3)
    ok = .false.
    if (.not.gpu_enabled .or. .not.c_associated(h)) return
    nn = int(n,c_int)
    ierr = c2gpu_promol_grid(h,nn,f)
    call gpu_check("promolecular_grid",ierr)
    ok = (ierr == 0)
  end subroutine gpu_promolecular_grid

  !> mesh%fill (meshmod@proc.f90:446-470) on the GPU for im_rho / im_gradrho / im_gkin: one call per property column.
  !> iprop: 0 rho (res%f), 1 gradrho (res%gfmod), 2 gkin (res%gkin; wavefunction fields only).
  subroutine gpu_fillmesh(h,x,iprop,periodic,fcol,ok)
    type(c_ptr), intent(in) :: h
    real*8, intent(in) :: x(:,:)          !< m%x(3,n)
    integer, intent(in) :: iprop
    logical, intent(in) :: periodic
    real*8, intent(inout) :: fcol(:)      !< m%f(:,j)
    logical, intent(out) :: ok
    real(c_double), allocatable :: res(:,:)
    integer(c_int) :: ierr, iper
    integer(c_long) :: n
    ok = .false.
    if (.not.c_associated(h)) return
    n = size(x,2)
    iper = merge(1_c_int,0_c_int,periodic)
    if (iprop == 2) then
       ierr = c2gpu_eval_points_gkin(h,n,x,iper,fcol)
    else
       allocate(res(16,n))
       ierr = c2gpu_eval_points(h,n,x,int(iprop,c_int),iper,res)
       if (ierr == 0) fcol = res(merge(1,14,iprop==0),:)      ! res%f, res%gfmod
    end if
    if (ierr == -1) return
    call gpu_check("gpu_fillmesh",ierr)
    ok = .true.
  end subroutine gpu_fillmesh

  !> mesh%fill with prop > 100 (MO values, grd nder=-1 -> calculate_mo): hmo is the one-MO handle made by
  !> gpu_wfn_handle over the row cmo(imo,:) (see INTEGRATION.md).
  subroutine gpu_fillmesh_mo(hmo,x,periodic,fcol,ok)
    type(c_ptr), intent(in) :: hmo
    real*8, intent(in) :: x(:,:)
    logical, intent(in) :: periodic
    real*8, intent(inout) :: fcol(:)
    logical, intent(out) :: ok
    integer(c_int) :: ierr
    ok = .false.
    if (.not.c_associated(hmo)) return
    ierr = c2gpu_eval_points_mo(hmo,int(size(x,2),c_long),x,merge(1_c_int,0_c_int,periodic),fcol)
    if (ierr == -1) return
    call gpu_check("gpu_fillmesh_mo",ierr)
    ok = .true.
  end subroutine gpu_fillmesh_mo

end module gpu_interface
