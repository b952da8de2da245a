!> ISO_C_BINDING shim: the Fortran side of the drop-in boundary (include/omg_gpu.h).
!>
!> A host program keeps `use m_octree_mg` and the reference's own setup modules (m_data_structures,
!> m_build_tree, m_load_balance, m_allocate_storage); this module replaces the BODIES of the cycle entry
!> points of src/m_multigrid.f90 (mg_fas_fmg :84-147, mg_fas_vcycle :150-243, mg_apply_op :439-456) and the
!> level-wide entry points of src/m_ghost_cells.f90 (:120-175), src/m_restrict.f90 (:72-114) and
!> src/m_prolong.f90 (:51-85) by calls into libomg_b200.so.  There is no CPU fallback: a failing call
!> `error stop`s with the library's message, exactly like the reference's own error handling.
!>
!> This file cannot be compiled in the build image (no Fortran compiler, SURVEY.md F2); the C++ mirror in
!> octree-mg_b200/host/ makes the same calls in the same order and is what the tests drive.  Build where a
!> compiler exists with:   mpif90 -cpp -DNDIM=3 -I<reference>/lib_3d -c m_gpu_multigrid.f90
!> and link  -L<repo>/octree-mg_b200 -lomg_b200  (see INTEGRATION.md).
module m_gpu_multigrid
  use, intrinsic :: iso_c_binding
  use m_data_structures

  implicit none
  private

  integer, parameter :: n_lvl = mg_lvl_hi - mg_lvl_lo + 1

  !> Mirror of `struct omg_tree_desc` (include/omg_gpu.h)
  type, bind(C) :: omg_tree_desc
     integer(c_int32_t) :: n_boxes, box_size, lowest_lvl, highest_lvl, first_normal_lvl
     integer(c_int32_t) :: my_rank, n_cpu, n_extra_vars
     integer(c_int32_t) :: periodic(3)
     type(c_ptr)        :: box_size_lvl, dr, box_rank, box_lvl, box_ix, box_parent
     type(c_ptr)        :: box_children, box_neighbors, lvl_ptr, lvl_ids
  end type omg_tree_desc

  !> Mirror of `struct omg_params`
  type, bind(C) :: omg_params
     integer(c_int32_t) :: operator_type, smoother_type, geometry_type
     integer(c_int32_t) :: n_cycle_down, n_cycle_up, max_coarse_cycles, subtract_mean
     real(c_double)     :: residual_coarse_abs, residual_coarse_rel, lambda
  end type omg_params

  interface
     integer(c_int) function omg_create(device, handle) bind(C, name="omg_create")
       import; integer(c_int), value :: device; type(c_ptr) :: handle
     end function omg_create
     integer(c_int) function omg_destroy(h) bind(C, name="omg_destroy")
       import; type(c_ptr), value :: h
     end function omg_destroy
     type(c_ptr) function omg_last_error(h) bind(C, name="omg_last_error")
       import; type(c_ptr), value :: h
     end function omg_last_error
     integer(c_int) function omg_set_tree(h, t) bind(C, name="omg_set_tree")
       import; type(c_ptr), value :: h; type(omg_tree_desc) :: t
     end function omg_set_tree
     integer(c_int) function omg_set_params(h, p) bind(C, name="omg_set_params")
       import; type(c_ptr), value :: h; type(omg_params) :: p
     end function omg_set_params
     integer(c_int) function omg_set_bc_const(h, iv, nb, bc_type, bc_value) bind(C, name="omg_set_bc_const")
       import; type(c_ptr), value :: h; integer(c_int), value :: iv, nb, bc_type; real(c_double), value :: bc_value
     end function omg_set_bc_const
     integer(c_int) function omg_set_bc_plane(h, id, nb, bc_type, plane) bind(C, name="omg_set_bc_plane")
       import; type(c_ptr), value :: h; integer(c_int), value :: id, nb, bc_type; real(c_double) :: plane(*)
     end function omg_set_bc_plane
     integer(c_int) function omg_push_var(h, iv, n, ids, ptrs) bind(C, name="omg_push_var")
       import; type(c_ptr), value :: h; integer(c_int), value :: iv, n; integer(c_int32_t) :: ids(*); type(c_ptr) :: ptrs(*)
     end function omg_push_var
     integer(c_int) function omg_pull_var(h, iv, n, ids, ptrs, with_ghosts) bind(C, name="omg_pull_var")
       import; type(c_ptr), value :: h; integer(c_int), value :: iv, n, with_ghosts
       integer(c_int32_t) :: ids(*); type(c_ptr) :: ptrs(*)
     end function omg_pull_var
     integer(c_int) function omg_fas_vcycle(h, has_highest, highest_lvl, standalone, want_res, max_res) &
          bind(C, name="omg_fas_vcycle")
       import; type(c_ptr), value :: h; integer(c_int), value :: has_highest, highest_lvl, standalone, want_res
       real(c_double) :: max_res
     end function omg_fas_vcycle
     integer(c_int) function omg_fas_fmg(h, have_guess, want_res, max_res) bind(C, name="omg_fas_fmg")
       import; type(c_ptr), value :: h; integer(c_int), value :: have_guess, want_res; real(c_double) :: max_res
     end function omg_fas_fmg
     integer(c_int) function omg_fill_ghost_cells_lvl(h, lvl, iv) bind(C, name="omg_fill_ghost_cells_lvl")
       import; type(c_ptr), value :: h; integer(c_int), value :: lvl, iv
     end function omg_fill_ghost_cells_lvl
     integer(c_int) function omg_restrict_lvl(h, iv, lvl) bind(C, name="omg_restrict_lvl")
       import; type(c_ptr), value :: h; integer(c_int), value :: iv, lvl
     end function omg_restrict_lvl
     integer(c_int) function omg_prolong(h, lvl, iv, iv_to, add) bind(C, name="omg_prolong")
       import; type(c_ptr), value :: h; integer(c_int), value :: lvl, iv, iv_to, add
     end function omg_prolong
     integer(c_int) function omg_apply_op(h, i_out) bind(C, name="omg_apply_op")
       import; type(c_ptr), value :: h; integer(c_int), value :: i_out
     end function omg_apply_op
     integer(c_int) function omg_diffusion_set_rhs(h, f1, f2) bind(C, name="omg_diffusion_set_rhs")
       import; type(c_ptr), value :: h; real(c_double), value :: f1, f2
     end function omg_diffusion_set_rhs
     integer(c_int) function omg_copy_to_tree(h, iv, lvl_lo, lvl_hi, leaves_only, with_ghosts, buf, n) &
          bind(C, name="omg_copy_to_tree")
       import; type(c_ptr), value :: h; integer(c_int), value :: iv, lvl_lo, lvl_hi, leaves_only, with_ghosts
       real(c_double), intent(in) :: buf(*); integer(c_int64_t), value :: n
     end function omg_copy_to_tree
     integer(c_int) function omg_copy_from_tree(h, iv, lvl_lo, lvl_hi, leaves_only, with_ghosts, buf, n) &
          bind(C, name="omg_copy_from_tree")
       import; type(c_ptr), value :: h; integer(c_int), value :: iv, lvl_lo, lvl_hi, leaves_only, with_ghosts
       real(c_double), intent(inout) :: buf(*); integer(c_int64_t), value :: n
     end function omg_copy_from_tree
     integer(c_int64_t) function omg_copy_size(h, lvl_lo, lvl_hi, leaves_only, with_ghosts) bind(C, name="omg_copy_size")
       import; type(c_ptr), value :: h; integer(c_int), value :: lvl_lo, lvl_hi, leaves_only, with_ghosts
     end function omg_copy_size
     integer(c_int) function omg_update_tree(h, t) bind(C, name="omg_update_tree")
       import; type(c_ptr), value :: h; type(omg_tree_desc), intent(in) :: t
     end function omg_update_tree
     integer(c_int) function omg_free_space_setup(h, fft_lvl, r_min, n_kernel, dr) bind(C, name="omg_free_space_setup")
       import; type(c_ptr), value :: h; integer(c_int), value :: fft_lvl
       real(c_double), intent(in) :: r_min(3), dr(3); integer(c_int), intent(in) :: n_kernel(3)
     end function omg_free_space_setup
     integer(c_int) function omg_free_space_solve(h, rhs_fac) bind(C, name="omg_free_space_solve")
       import; type(c_ptr), value :: h; real(c_double), value :: rhs_fac
     end function omg_free_space_solve
     integer(c_int) function omg_get_bc_plane(h, id, nb, bc_type, plane) bind(C, name="omg_get_bc_plane")
       import; type(c_ptr), value :: h; integer(c_int), value :: id, nb
       integer(c_int), intent(out) :: bc_type; real(c_double), intent(out) :: plane(*)
     end function omg_get_bc_plane
     integer(c_int) function omg_nccl_unique_id(uid) bind(C, name="omg_nccl_unique_id")
       import; character(kind=c_char) :: uid(128)
     end function omg_nccl_unique_id
     integer(c_int) function omg_comm_init(h, uid, rank, nranks) bind(C, name="omg_comm_init")
       import; type(c_ptr), value :: h; character(kind=c_char) :: uid(128); integer(c_int), value :: rank, nranks
     end function omg_comm_init
     function c_strlen(s) bind(C, name="strlen") result(n)
       import; type(c_ptr), value :: s; integer(c_size_t) :: n
     end function c_strlen
  end interface

  !> Device handle and coherence mode of one mg_t (the reference type is not extended: the shim keeps
  !> its state beside it; one tree per process)
  type :: gpu_state_t
     type(c_ptr) :: handle = c_null_ptr
     logical     :: resident = .false.   !< .true.: the host brackets its own cc accesses with mg_gpu_push/pull
     real(dp)    :: lambda = 0.0_dp      !< helmholtz / vhelmholtz / ahelmholtz lambda (module globals in the reference)
     logical     :: free_initialized = .false.   !< type(mg_free_bc_t) of src/m_free_space.f90:9-22: what the host keeps of it
     integer     :: free_fft_lvl = -1
  end type gpu_state_t

  type(gpu_state_t), save :: gpu   !< one tree per process, like the reference's module-level timers

  public :: mg_gpu_attach, mg_gpu_detach, mg_gpu_set_resident, mg_gpu_set_lambda
  public :: mg_gpu_push, mg_gpu_pull
  public :: mg_fas_fmg, mg_fas_vcycle, mg_apply_op
  public :: diffusion_solve, diffusion_solve_vcoeff, diffusion_solve_acoeff
  public :: mg_gpu_copy_size, mg_gpu_copy_to_tree, mg_gpu_copy_from_tree
  public :: mg_fill_ghost_cells_lvl, mg_fill_ghost_cells, mg_restrict_lvl, mg_restrict, mg_prolong
  public :: mg_gpu_update_tree, mg_poisson_free_3d

contains

  subroutine check(rc, where)
    integer(c_int), intent(in)   :: rc
    character(len=*), intent(in) :: where
    character(kind=c_char), pointer :: msg(:)
    type(c_ptr) :: p
    integer     :: n
    if (rc == 0) return
    p = omg_last_error(gpu%handle)
    n = int(c_strlen(p))
    call c_f_pointer(p, msg, [n])
    print *, where, ": ", msg(1:n)
    error stop "octree-mg GPU path failed (there is no CPU fallback)"
  end subroutine check

  !> Call after mg_allocate_storage (src/m_allocate_storage.f90:51): flattens the tree metadata of mg_t
  !> into omg_tree_desc, creates the device mirror on `device` and, with several ranks, the NCCL communicator
  !> (the unique id is broadcast over mg%comm -- the only MPI call the GPU path keeps).
  subroutine mg_gpu_attach(mg, device)
    type(mg_t), intent(inout), target :: mg
    integer, intent(in)               :: device
    call hand_over_tree(mg, device, .false.)
  end subroutine mg_gpu_attach

  !> On-device regrid / re-partition: the host has changed the tree of an attached mg_t (mg_add_children +
  !> mg_set_neighbors_lvl + mg_set_refinement_boundaries, src/m_build_tree.f90:296-367; mg_load_balance,
  !> src/m_load_balance.f90:71-136; cc arrays of new boxes allocated by the caller).  Boxes that exist before and after
  !> keep their device data (omg_update_tree); new boxes start from zero.  In drop-in mode the host arrays are pushed
  !> before and pulled after, so that boxes which change owner take all their variables along.
  subroutine mg_gpu_update_tree(mg)
    type(mg_t), intent(inout), target :: mg
    integer :: iv
    if (.not. c_associated(gpu%handle)) error stop "mg_gpu_update_tree: call mg_gpu_attach first"
    call hand_over_tree(mg, 0, .true.)
    gpu%free_initialized = .false.
    if (.not. gpu%resident) then
       call push_params(mg)
       do iv = 1, mg_num_vars + mg%n_extra_vars
          call mg_gpu_pull(mg, iv, iv == mg_iphi .or. iv >= mg_iveps)
       end do
    end if
  end subroutine mg_gpu_update_tree

  subroutine hand_over_tree(mg, device, update)
    use mpi
    type(mg_t), intent(inout), target :: mg
    integer, intent(in)               :: device
    logical, intent(in)               :: update
    type(omg_tree_desc)               :: t
    integer(c_int32_t), allocatable, target :: rank(:), lvl(:), ix(:, :), parent(:), children(:, :)
    integer(c_int32_t), allocatable, target :: neighbors(:, :), lvl_ptr(:), lvl_ids(:), bsl(:)
    real(c_double), allocatable, target     :: dr(:, :)
    character(kind=c_char)                  :: uid(128)
    integer :: id, l, n, n_ids, ierr

    if (.not. mg%is_allocated) error stop "mg_gpu_attach: call mg_allocate_storage first"
    n = mg%n_boxes
    allocate(rank(n), lvl(n), ix(3, n), parent(n), children(8, n), neighbors(6, n))
    do id = 1, n
       rank(id)         = mg%boxes(id)%rank
       lvl(id)          = mg%boxes(id)%lvl
       ix(:, id)        = mg%boxes(id)%ix
       parent(id)       = mg%boxes(id)%parent
       children(:, id)  = mg%boxes(id)%children
       neighbors(:, id) = mg%boxes(id)%neighbors
    end do

    n_ids = 0
    do l = mg%lowest_lvl, mg%highest_lvl
       n_ids = n_ids + size(mg%lvls(l)%ids)
    end do
    allocate(lvl_ptr(n_lvl + 1), lvl_ids(n_ids), bsl(n_lvl), dr(3, n_lvl))
    lvl_ptr(1) = 0
    n_ids = 0
    do l = mg_lvl_lo, mg_lvl_hi
       if (l >= mg%lowest_lvl .and. l <= mg%highest_lvl) then
          lvl_ids(n_ids+1:n_ids+size(mg%lvls(l)%ids)) = mg%lvls(l)%ids
          n_ids = n_ids + size(mg%lvls(l)%ids)
       end if
       lvl_ptr(l - mg_lvl_lo + 2) = n_ids
       bsl(l - mg_lvl_lo + 1)     = mg%box_size_lvl(l)
       dr(:, l - mg_lvl_lo + 1)   = mg%dr(:, l)
    end do

    t%n_boxes = n;                    t%box_size = mg%box_size
    t%lowest_lvl = mg%lowest_lvl;     t%highest_lvl = mg%highest_lvl
    t%first_normal_lvl = mg%first_normal_lvl
    t%my_rank = mg%my_rank;           t%n_cpu = mg%n_cpu
    t%n_extra_vars = mg%n_extra_vars
    t%periodic = merge(1, 0, mg%periodic)
    t%box_size_lvl = c_loc(bsl);      t%dr = c_loc(dr)
    t%box_rank = c_loc(rank);         t%box_lvl = c_loc(lvl);        t%box_ix = c_loc(ix)
    t%box_parent = c_loc(parent);     t%box_children = c_loc(children)
    t%box_neighbors = c_loc(neighbors)
    t%lvl_ptr = c_loc(lvl_ptr);       t%lvl_ids = c_loc(lvl_ids)

    if (.not. c_associated(gpu%handle)) then
       if (omg_create(int(device, c_int), gpu%handle) /= 0) &
            error stop "mg_gpu_attach: no CUDA device available (there is no CPU fallback)"
    end if
    if (update) then
       call check(omg_update_tree(gpu%handle, t), "omg_update_tree")   ! keeps the communicator, renews the pool mappings
       return
    end if
    call check(omg_set_tree(gpu%handle, t), "omg_set_tree")

    if (mg%n_cpu > 1) then
       if (mg%my_rank == 0) call check(omg_nccl_unique_id(uid), "omg_nccl_unique_id")
       call mpi_bcast(uid, 128, MPI_CHARACTER, 0, mg%comm, ierr)
       call check(omg_comm_init(gpu%handle, uid, int(mg%my_rank, c_int), int(mg%n_cpu, c_int)), "omg_comm_init")
    end if
  end subroutine hand_over_tree

  !> Call from / instead of mg_deallocate_storage (src/m_allocate_storage.f90:14-47)
  subroutine mg_gpu_detach()
    integer(c_int) :: rc
    if (c_associated(gpu%handle)) rc = omg_destroy(gpu%handle)
    gpu%handle = c_null_ptr
  end subroutine mg_gpu_detach

  subroutine mg_gpu_set_resident(resident)
    logical, intent(in) :: resident
    gpu%resident = resident
  end subroutine mg_gpu_set_resident

  !> helmholtz_set_lambda / vhelmholtz_set_lambda / ahelmholtz_set_lambda (src/m_helmholtz.f90:39-46)
  subroutine mg_gpu_set_lambda(lambda)
    real(dp), intent(in) :: lambda
    if (lambda < 0) error stop "helmholtz_set_lambda: lambda < 0 not allowed"
    gpu%lambda = lambda
  end subroutine mg_gpu_set_lambda

  !> mg_t scalars, boundary-condition constants and the planes of boundary_cond callbacks
  subroutine push_params(mg)
    type(mg_t), intent(inout) :: mg
    type(omg_params)          :: p
    integer                   :: iv, nb, lvl, i, id, nc, bc_type
    real(dp), allocatable     :: bc(:, :)

    p%operator_type = mg%operator_type;   p%smoother_type = mg%smoother_type
    p%geometry_type = mg%geometry_type
    p%n_cycle_down = mg%n_cycle_down;     p%n_cycle_up = mg%n_cycle_up
    p%max_coarse_cycles = mg%max_coarse_cycles
    p%subtract_mean = merge(1, 0, mg%subtract_mean)
    p%residual_coarse_abs = mg%residual_coarse_abs
    p%residual_coarse_rel = mg%residual_coarse_rel
    p%lambda = gpu%lambda
    call check(omg_set_params(gpu%handle, p), "omg_set_params")

    do iv = 1, mg_num_vars + mg%n_extra_vars
       do nb = 1, mg_num_neighbors
          call check(omg_set_bc_const(gpu%handle, iv, nb, mg%bc(nb, iv)%bc_type, mg%bc(nb, iv)%bc_value), &
               "omg_set_bc_const")
       end do
    end do

    ! What set_ghost_cells (src/m_ghost_cells.f90:262-279) would obtain on every fill: the planes kept by
    ! mg_phi_bc_store in the rhs ghost cells, or the callback's values; constants need nothing more.
    do lvl = mg%lowest_lvl, mg%highest_lvl
       nc = mg%box_size_lvl(lvl)
       allocate(bc(nc, nc))
       do i = 1, size(mg%lvls(lvl)%my_ids)
          id = mg%lvls(lvl)%my_ids(i)
          do nb = 1, mg_num_neighbors
             if (mg%boxes(id)%neighbors(nb) >= mg_no_box) cycle
             if (mg%phi_bc_data_stored) then
                call box_get_face(mg%boxes(id), nb, nc, mg_irhs, bc)
                bc_type = mg%boxes(id)%neighbors(nb)
             else if (associated(mg%bc(nb, mg_iphi)%boundary_cond)) then
                call mg%bc(nb, mg_iphi)%boundary_cond(mg%boxes(id), nc, mg_iphi, nb, bc_type, bc)
             else
                cycle
             end if
             call check(omg_set_bc_plane(gpu%handle, id, nb, bc_type, bc), "omg_set_bc_plane")
          end do
       end do
       deallocate(bc)
    end do
  end subroutine push_params

  !> Ghost plane (0 or nc+1) of variable iv on side nb, tangential range 1:nc
  subroutine box_get_face(box, nb, nc, iv, plane)
    type(mg_box_t), intent(in) :: box
    integer, intent(in)        :: nb, nc, iv
    real(dp), intent(out)      :: plane(nc, nc)
    integer                    :: g
    g = merge(0, nc + 1, mg_neighb_low(nb))
    select case (mg_neighb_dim(nb))
    case (1); plane = box%cc(g, 1:nc, 1:nc, iv)
    case (2); plane = box%cc(1:nc, g, 1:nc, iv)
    case (3); plane = box%cc(1:nc, 1:nc, g, iv)
    end select
  end subroutine box_get_face

  !> Host -> device copy of variable iv of all owned boxes (interiors; ghost faces too for iv >= 5)
  subroutine mg_gpu_push(mg, iv)
    type(mg_t), intent(inout), target :: mg
    integer, intent(in)               :: iv
    call transfer(mg, iv, mg%lowest_lvl, mg%highest_lvl, .true., .false.)
  end subroutine mg_gpu_push

  !> Device -> host copy; with_ghosts also writes the six ghost faces as the reference would hold them
  subroutine mg_gpu_pull(mg, iv, with_ghosts)
    type(mg_t), intent(inout), target :: mg
    integer, intent(in)               :: iv
    logical, intent(in)               :: with_ghosts
    call transfer(mg, iv, mg%lowest_lvl, mg%highest_lvl, .false., with_ghosts)
  end subroutine mg_gpu_pull

  subroutine transfer(mg, iv, lvl_lo, lvl_hi, push, with_ghosts)
    type(mg_t), intent(inout), target :: mg
    integer, intent(in)               :: iv, lvl_lo, lvl_hi
    logical, intent(in)               :: push, with_ghosts
    integer(c_int32_t), allocatable   :: ids(:)
    type(c_ptr), allocatable          :: ptrs(:)
    integer                           :: lvl, i, n, id

    n = 0
    do lvl = max(lvl_lo, mg%lowest_lvl), min(lvl_hi, mg%highest_lvl)
       n = n + size(mg%lvls(lvl)%my_ids)
    end do
    allocate(ids(n), ptrs(n))
    n = 0
    do lvl = max(lvl_lo, mg%lowest_lvl), min(lvl_hi, mg%highest_lvl)
       do i = 1, size(mg%lvls(lvl)%my_ids)
          id = mg%lvls(lvl)%my_ids(i)
          n = n + 1
          ids(n)  = id
          ptrs(n) = c_loc(mg%boxes(id)%cc(0, 0, 0, iv))   ! contiguous (nc+2)^3 slice of variable iv
       end do
    end do
    if (push) then
       call check(omg_push_var(gpu%handle, iv, n, ids, ptrs), "omg_push_var")
    else
       call check(omg_pull_var(gpu%handle, iv, n, ids, ptrs, merge(1, 0, with_ghosts)), "omg_pull_var")
    end if
  end subroutine transfer

  integer function n_eps_vars(mg)
    type(mg_t), intent(in) :: mg
    select case (mg%operator_type)
    case (mg_vlaplacian, mg_vhelmholtz); n_eps_vars = 1
    case (mg_ahelmholtz);                n_eps_vars = 3
    case default;                        n_eps_vars = 0
    end select
  end function n_eps_vars

  !> Strict drop-in coherence: what the cycle reads is pushed, what it writes is pulled (SURVEY.md 8b)
  subroutine sync_in(mg)
    type(mg_t), intent(inout) :: mg
    integer :: e
    call push_params(mg)
    if (gpu%resident) return
    call mg_gpu_push(mg, mg_iphi)
    call mg_gpu_push(mg, mg_irhs)
    do e = 1, n_eps_vars(mg)
       call mg_gpu_push(mg, mg_iveps + e - 1)
    end do
  end subroutine sync_in

  subroutine sync_out(mg)
    type(mg_t), intent(inout) :: mg
    if (gpu%resident) return
    call mg_gpu_pull(mg, mg_iphi, .true.)
    call mg_gpu_pull(mg, mg_irhs, .false.)
    call mg_gpu_pull(mg, mg_iold, .true.)
    call mg_gpu_pull(mg, mg_ires, .false.)
  end subroutine sync_out

  !> Same interface as src/m_multigrid.f90:84-87
  subroutine mg_fas_fmg(mg, have_guess, max_res)
    type(mg_t), intent(inout)       :: mg
    logical, intent(in)             :: have_guess
    real(dp), intent(out), optional :: max_res
    real(c_double)                  :: res
    call sync_in(mg)
    call check(omg_fas_fmg(gpu%handle, merge(1, 0, have_guess), merge(1, 0, present(max_res)), res), "mg_fas_fmg")
    if (present(max_res)) max_res = res
    call sync_out(mg)
  end subroutine mg_fas_fmg

  !> Same interface as src/m_multigrid.f90:150-156
  subroutine mg_fas_vcycle(mg, highest_lvl, max_res, standalone)
    type(mg_t), intent(inout)       :: mg
    integer, intent(in), optional   :: highest_lvl
    real(dp), intent(out), optional :: max_res
    logical, intent(in), optional   :: standalone
    real(c_double)                  :: res
    integer(c_int)                  :: has, top, alone
    has = merge(1, 0, present(highest_lvl)); top = 0; alone = 1
    if (present(highest_lvl)) top = highest_lvl
    if (present(standalone)) alone = merge(1, 0, standalone)
    call sync_in(mg)
    call check(omg_fas_vcycle(gpu%handle, has, top, alone, merge(1, 0, present(max_res)), res), "mg_fas_vcycle")
    if (present(max_res)) max_res = res
    call sync_out(mg)
  end subroutine mg_fas_vcycle

  !> src/m_multigrid.f90:439-442 with the built-in operator (a user-supplied `op` is not supported on the device)
  subroutine mg_apply_op(mg, i_out)
    type(mg_t), intent(inout) :: mg
    integer, intent(in)       :: i_out
    call sync_in(mg)
    call check(omg_apply_op(gpu%handle, i_out), "mg_apply_op")
    if (.not. gpu%resident) call mg_gpu_pull(mg, i_out, .false.)
  end subroutine mg_apply_op

  !> src/m_ghost_cells.f90:131-175
  subroutine mg_fill_ghost_cells_lvl(mg, lvl, iv)
    type(mg_t), intent(inout) :: mg
    integer, intent(in)       :: lvl, iv
    call push_params(mg)
    if (.not. gpu%resident) call transfer(mg, iv, lvl - 1, lvl, .true., .false.)
    call check(omg_fill_ghost_cells_lvl(gpu%handle, lvl, iv), "mg_fill_ghost_cells_lvl")
    if (.not. gpu%resident) call transfer(mg, iv, lvl, lvl, .false., .true.)
  end subroutine mg_fill_ghost_cells_lvl

  !> src/m_ghost_cells.f90:120-128
  subroutine mg_fill_ghost_cells(mg, iv)
    type(mg_t), intent(inout) :: mg
    integer, intent(in)       :: iv
    integer                   :: lvl
    do lvl = mg%lowest_lvl, mg%highest_lvl
       call mg_fill_ghost_cells_lvl(mg, lvl, iv)
    end do
  end subroutine mg_fill_ghost_cells

  !> src/m_restrict.f90:83-114
  subroutine mg_restrict_lvl(mg, iv, lvl)
    type(mg_t), intent(inout) :: mg
    integer, intent(in)       :: iv, lvl
    call push_params(mg)
    if (.not. gpu%resident) call transfer(mg, iv, lvl - 1, lvl, .true., .false.)
    call check(omg_restrict_lvl(gpu%handle, iv, lvl), "mg_restrict_lvl")
    if (.not. gpu%resident) call transfer(mg, iv, lvl - 1, lvl - 1, .false., .false.)
  end subroutine mg_restrict_lvl

  !> src/m_restrict.f90:72-80
  subroutine mg_restrict(mg, iv)
    type(mg_t), intent(inout) :: mg
    integer, intent(in)       :: iv
    integer                   :: lvl
    do lvl = mg%highest_lvl, mg%lowest_lvl + 1, -1
       call mg_restrict_lvl(mg, iv, lvl)
    end do
  end subroutine mg_restrict

  !> src/m_prolong.f90:51-58 with the default method mg_prolong_sparse (a user `method` is not supported)
  subroutine mg_prolong(mg, lvl, iv, iv_to, add)
    type(mg_t), intent(inout) :: mg
    integer, intent(in)       :: lvl, iv, iv_to
    logical, intent(in)       :: add
    call push_params(mg)
    if (.not. gpu%resident) then
       call transfer(mg, iv, lvl, lvl, .true., .false.)
       call transfer(mg, iv_to, lvl + 1, lvl + 1, .true., .false.)
    end if
    call check(omg_prolong(gpu%handle, lvl, iv, iv_to, merge(1, 0, add)), "mg_prolong")
    if (.not. gpu%resident) call transfer(mg, iv_to, lvl + 1, lvl + 1, .false., .false.)
  end subroutine mg_prolong

  !> The three solvers of src/m_diffusion.f90 (:19-58, :64-106, :113-157) share this body: operator `op`,
  !> lambda = order/scale, rhs from set_rhs (:159-173) on the device, FMG with a guess, V-cycles until max_res.
  !> One push on entry and one pull on exit; in between only the residual crosses PCIe.
  subroutine diffusion_common(mg, op, scale, order, max_res)
    use m_multigrid, only: mg_set_methods    ! stays the reference's (host-side bookkeeping only, :27-60)
    type(mg_t), intent(inout) :: mg
    integer, intent(in)       :: op, order
    real(dp), intent(in)      :: scale, max_res
    integer, parameter        :: max_its = 10
    integer                   :: n, e
    real(c_double)            :: res
    logical                   :: was_resident

    if (order /= 1 .and. order /= 2) error stop "diffusion_solve: order should be 1 or 2"
    mg%operator_type = op
    call mg_set_methods(mg)           ! the host's own (reference) routine: bc of eps vars, subtract_mean, checks
    was_resident = gpu%resident
    if (.not. was_resident) then
       call mg_gpu_push(mg, mg_iphi)
       call mg_gpu_push(mg, mg_irhs)
       do e = 1, n_eps_vars(mg)
          call mg_gpu_push(mg, mg_iveps + e - 1)
       end do
    end if
    if (order == 1) then
       call mg_gpu_set_lambda(1/scale)
       call push_params(mg)
       call check(omg_diffusion_set_rhs(gpu%handle, -1/scale, 0.0_dp), "diffusion_solve")
    else
       call mg_gpu_set_lambda(0.0_dp)
       call push_params(mg)
       call check(omg_apply_op(gpu%handle, mg_irhs), "diffusion_solve")
       call mg_gpu_set_lambda(2/scale)
       call push_params(mg)
       call check(omg_diffusion_set_rhs(gpu%handle, -2/scale, -1.0_dp), "diffusion_solve")
    end if
    call check(omg_fas_fmg(gpu%handle, 1, 1, res), "diffusion_solve")
    do n = 1, max_its
       if (res <= max_res) exit
       call check(omg_fas_vcycle(gpu%handle, 0, 0, 1, 1, res), "diffusion_solve")
    end do
    call sync_out(mg)
    if (n == max_its + 1) then
       if (mg%my_rank == 0) print *, "Did you specify boundary conditions correctly?"
       error stop "diffusion_solve: no convergence"
    end if
  end subroutine diffusion_common

  !> Same interface as src/m_diffusion.f90:19
  subroutine diffusion_solve(mg, dt, diffusion_coeff, order, max_res)
    type(mg_t), intent(inout) :: mg
    real(dp), intent(in)      :: dt, diffusion_coeff, max_res
    integer, intent(in)       :: order
    call diffusion_common(mg, mg_helmholtz, dt * diffusion_coeff, order, max_res)
  end subroutine diffusion_solve

  !> Same interface as src/m_diffusion.f90:64 (coefficient in mg_iveps on all levels)
  subroutine diffusion_solve_vcoeff(mg, dt, order, max_res)
    type(mg_t), intent(inout) :: mg
    real(dp), intent(in)      :: dt, max_res
    integer, intent(in)       :: order
    call diffusion_common(mg, mg_vhelmholtz, dt, order, max_res)
  end subroutine diffusion_solve_vcoeff

  !> Same interface as src/m_diffusion.f90:113 (coefficients in mg_iveps1..3 on all levels)
  subroutine diffusion_solve_acoeff(mg, dt, order, max_res)
    type(mg_t), intent(inout) :: mg
    real(dp), intent(in)      :: dt, max_res
    integer, intent(in)       :: order
    call diffusion_common(mg, mg_ahelmholtz, dt, order, max_res)
  end subroutine diffusion_solve_acoeff

  !> Same interface as src/m_free_space.f90:36-50.  The FFT level is chosen as there (:75-86); the coarse free-space
  !> solve, the Dirichlet planes of all levels and the initial guess are made on the device (omg_free_space_setup builds
  !> the kernel createKernel would build, omg_free_space_solve replaces :140-180), the rest is the reference's sequence
  !> of level-wide calls (:183-206).
  subroutine mg_poisson_free_3d(mg, new_rhs, max_fft_frac, fmgcycle, max_res)
    type(mg_t), intent(inout)       :: mg
    logical, intent(in)             :: new_rhs, fmgcycle
    real(dp), intent(in)            :: max_fft_frac
    real(dp), intent(out), optional :: max_res
    real(dp), parameter             :: rhs_fac = -1 / (4 * acos(-1.0_dp))
    integer                         :: lvl, fft_lvl, n, nb, id, nc, bc_type
    integer(c_int)                  :: nk(3)
    integer(i8)                     :: n_lvl, n_total
    logical                         :: new_fft_grid
    real(c_double)                  :: res
    real(dp), allocatable           :: plane(:, :)

    if (.not. gpu%free_initialized .and. .not. new_rhs) &
         error stop "mg_poisson_free_3d: first call requires new_rhs = .true."
    if (mg%geometry_type /= mg_cartesian) error stop "mg_poisson_free_3d: Cartesian 3D geometry required"
    if (mg%operator_type /= mg_laplacian) error stop "mg_poisson_free_3d: laplacian operator required"

    n_total = mg_number_of_unknowns(mg)
    do lvl = mg_highest_uniform_lvl(mg), mg%lowest_lvl+1, -1
       n_lvl = size(mg%lvls(lvl)%ids) * int(mg%box_size**3, i8)
       if (n_lvl <= max_fft_frac * n_total) exit
    end do
    fft_lvl      = lvl
    new_fft_grid = (.not. gpu%free_initialized .or. gpu%free_fft_lvl /= fft_lvl)

    ! the boundary planes live on the device from here on
    do n = 1, mg_num_neighbors
       nullify(mg%bc(n, mg_iphi)%boundary_cond)
       mg%bc(n, mg_iphi)%bc_type = mg_bc_dirichlet
    end do
    call sync_in(mg)

    if (new_fft_grid) then
       gpu%free_initialized = .false.
       gpu%free_fft_lvl     = fft_lvl
       do lvl = mg%highest_lvl, fft_lvl+1, -1
          call check(omg_restrict_lvl(gpu%handle, mg_irhs, lvl), "mg_poisson_free_3d")
       end do
       ! createKernel('F', nx(1), nx(3), nx(3), ...) of the reference (:119), nx = domain size + 2
       nk = [mg%domain_size_lvl(1, fft_lvl), mg%domain_size_lvl(3, fft_lvl), mg%domain_size_lvl(3, fft_lvl)] + 2
       call check(omg_free_space_setup(gpu%handle, fft_lvl, mg%r_min, nk, mg%dr(:, fft_lvl)), "mg_poisson_free_3d")
    end if

    if (new_rhs) then
       call check(omg_free_space_solve(gpu%handle, rhs_fac), "mg_poisson_free_3d")
       ! what mg_phi_bc_store (:170) leaves behind: the face types, and in drop-in mode the planes in the ghost cells of mg_irhs
       do lvl = mg%lowest_lvl, mg%highest_lvl
          nc = mg%box_size_lvl(lvl)
          if (allocated(plane)) deallocate(plane)
          allocate(plane(nc, nc))
          do n = 1, size(mg%lvls(lvl)%ids)
             id = mg%lvls(lvl)%ids(n)
             do nb = 1, mg_num_neighbors
                if (mg%boxes(id)%neighbors(nb) >= mg_no_box) cycle
                mg%boxes(id)%neighbors(nb) = mg_bc_dirichlet
                if (.not. gpu%resident .and. mg%boxes(id)%rank == mg%my_rank) then
                   call check(omg_get_bc_plane(gpu%handle, id, nb, bc_type, plane), "mg_poisson_free_3d")
                   call box_set_face(mg%boxes(id), nb, nc, mg_irhs, plane)
                end if
             end do
          end do
       end do
       mg%phi_bc_data_stored = .true.
       do lvl = fft_lvl, mg%lowest_lvl+1, -1
          call check(omg_restrict_lvl(gpu%handle, mg_iphi, lvl), "mg_poisson_free_3d")
       end do
       do lvl = fft_lvl, mg%highest_lvl-1
          call check(omg_prolong(gpu%handle, lvl, mg_iphi, mg_iphi, 0), "mg_poisson_free_3d")
          call check(omg_fill_ghost_cells_lvl(gpu%handle, lvl+1, mg_iphi), "mg_poisson_free_3d")
       end do
       gpu%free_initialized = .true.
    end if

    if (fft_lvl < mg%highest_lvl) then
       if (fmgcycle) then
          call check(omg_fas_fmg(gpu%handle, 1, merge(1, 0, present(max_res)), res), "mg_poisson_free_3d")
       else
          call check(omg_fas_vcycle(gpu%handle, 0, 0, 1, merge(1, 0, present(max_res)), res), "mg_poisson_free_3d")
       end if
       if (present(max_res)) max_res = res
    else if (present(max_res)) then
       max_res = 0.0_dp
    end if
    call sync_out(mg)
  end subroutine mg_poisson_free_3d

  !> box_set_gc of src/m_ghost_cells.f90:622-663 for one face
  subroutine box_set_face(box, nb, nc, iv, plane)
    type(mg_box_t), intent(inout) :: box
    integer, intent(in)           :: nb, nc, iv
    real(dp), intent(in)          :: plane(nc, nc)
    integer                       :: g
    g = merge(0, nc + 1, mg_neighb_low(nb))
    select case (mg_neighb_dim(nb))
    case (1); box%cc(g, 1:nc, 1:nc, iv) = plane
    case (2); box%cc(1:nc, g, 1:nc, iv) = plane
    case (3); box%cc(1:nc, 1:nc, g, iv) = plane
    end select
  end subroutine box_set_face

  !> Doubles a buffer of the two routines below holds: owned leaves (or all owned boxes), nc**3 or (nc+2)**3 per box
  integer(c_int64_t) function mg_gpu_copy_size(leaves_only, with_ghosts)
    logical, intent(in) :: leaves_only, with_ghosts
    mg_gpu_copy_size = omg_copy_size(gpu%handle, mg_lvl_lo, mg_lvl_hi, merge(1, 0, leaves_only), merge(1, 0, with_ghosts))
  end function mg_gpu_copy_size

  !> mg_copy_to_tree of the application coupling (coupling_amrvac/mod_multigrid_coupling.t:133-171) straight onto the
  !> device: buf = the grid leaves in my_leaves order, interior cells (or with one ghost layer), i fastest
  subroutine mg_gpu_copy_to_tree(mg, iv, buf, leaves_only, with_ghosts)
    type(mg_t), intent(inout) :: mg
    integer, intent(in)       :: iv
    real(dp), intent(in)      :: buf(:)
    logical, intent(in)       :: leaves_only, with_ghosts
    if (.not. mg%is_allocated) error stop "mg_copy_to_tree: tree not allocated yet"
    call push_params(mg)
    call check(omg_copy_to_tree(gpu%handle, iv, mg_lvl_lo, mg_lvl_hi, merge(1, 0, leaves_only), &
         merge(1, 0, with_ghosts), buf, int(size(buf), c_int64_t)), "mg_copy_to_tree")
  end subroutine mg_gpu_copy_to_tree

  !> mg_copy_from_tree / mg_copy_from_tree_gc (:206-267)
  subroutine mg_gpu_copy_from_tree(mg, iv, buf, leaves_only, with_ghosts)
    type(mg_t), intent(inout) :: mg
    integer, intent(in)       :: iv
    real(dp), intent(inout)   :: buf(:)
    logical, intent(in)       :: leaves_only, with_ghosts
    if (.not. mg%is_allocated) error stop "mg_copy_from_tree: tree not allocated yet"
    call push_params(mg)
    call check(omg_copy_from_tree(gpu%handle, iv, mg_lvl_lo, mg_lvl_hi, merge(1, 0, leaves_only), &
         merge(1, 0, with_ghosts), buf, int(size(buf), c_int64_t)), "mg_copy_from_tree")
  end subroutine mg_gpu_copy_from_tree

end module m_gpu_multigrid
