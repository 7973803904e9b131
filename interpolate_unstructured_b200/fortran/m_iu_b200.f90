!> m_iu_b200 -- ISO_C_BINDING shim over libiu_b200.so (include/iu_b200.h).
!>
!> Drop-in for the hot path of m_interp_unstructured (m_interp_unstructured.f90:77-101):
!> a user code replaces `use m_interp_unstructured` by `use m_iu_b200` and keeps calling
!> iu_read_grid, iu_get_cell, iu_get_cell_scalar_at, iu_get_icell_scalar_at,
!> iu_interpolate_at, iu_interpolate_scalar_at, iu_get_cell_through_neighbors and the
!> iu_get_*_data_index / iu_add_*_data bookkeeping with the reference's argument lists.
!> Scalar calls are batches of one query; the *_batch procedures are the new batched
!> variants (r(3, n), i_cell(n), var_at_r(n_vars, n)) that the GPU path is made for.
!>
!> The grid lives on the device: iu_grid_t here carries the opaque handle, the scalars of the
!> reference type (:25-35) and host mirrors of the three data tables, because callers of the
!> reference write ug%point_data directly (test_tetra.f90:38-40, test_trace_field.f90:20-23).
!> After such a write call iu_sync_point_data / iu_sync_cell_data / iu_sync_icell_data.
!>
!> Fatal conditions `error stop` with the library's message, like the reference.
!> NOTE: this image has no Fortran compiler, so this file is shipped uncompiled; the same
!> bindings are exercised through the ctypes mirror (interpolate_unstructured_b200/api.py)
!> and tests/test_abi_cpu.py checks that every symbol bound here is exported.
module m_iu_b200
  use iso_c_binding
  use iso_fortran_env, only: error_unit, int64

  implicit none
  private

  integer, parameter :: dp = kind(0.0d0)

  integer, parameter :: iu_triangle = 1
  integer, parameter :: iu_quad = 2
  integer, parameter :: iu_tetra = 3
  integer, parameter :: iu_ndim_cell_type(3) = [2, 2, 3]

  !> Device-resident counterpart of type(iu_grid_t), m_interp_unstructured.f90:24-59
  type iu_grid_t
     type(c_ptr) :: handle = c_null_ptr
     integer :: n_cells = 0
     integer :: n_points = 0
     integer :: n_points_per_cell = 0
     integer :: n_faces_per_cell = 0
     integer :: cell_type = 0
     integer :: n_point_data = 0
     integer :: n_cell_data = 0
     integer :: n_icell_data = 0
     real(dp) :: rmin(3) = 0.0_dp
     real(dp) :: rmax(3) = 0.0_dp
     !> host mirrors (written by the caller, pushed with iu_sync_*_data)
     real(dp), allocatable :: point_data(:, :)
     real(dp), allocatable :: cell_data(:, :)
     integer, allocatable  :: icell_data(:, :)
  end type iu_grid_t

  interface
     function c_last_error() bind(C, name="iu_b200_last_error") result(msg)
       import
       type(c_ptr) :: msg
     end function c_last_error

     function c_grid_from_binda(filename, scale, device, grid) bind(C, name="iu_b200_grid_from_binda") result(rc)
       import
       character(kind=c_char), intent(in) :: filename(*)
       real(c_double), value :: scale
       integer(c_int), value :: device
       type(c_ptr), intent(out) :: grid
       integer(c_int) :: rc
     end function c_grid_from_binda

     function c_grid_replicate(grid, n_devices, devices) bind(C, name="iu_b200_grid_replicate") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int), value :: n_devices
       integer(c_int), intent(in) :: devices(*)
       integer(c_int) :: rc
     end function c_grid_replicate

     function c_device_count(grid) bind(C, name="iu_b200_device_count") result(n)
       import
       type(c_ptr), value :: grid
       integer(c_int) :: n
     end function c_device_count

     function c_export_geometry(grid, cell_points, normals, volume, centers) &
          bind(C, name="iu_b200_export_geometry") result(rc)
       import
       type(c_ptr), value :: grid
       type(c_ptr), value :: cell_points, normals, volume
       real(c_double), intent(out) :: centers(3, *)
       integer(c_int) :: rc
     end function c_export_geometry

     function c_grid_destroy(grid) bind(C, name="iu_b200_grid_destroy") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int) :: rc
     end function c_grid_destroy

     function c_write_vtk(grid, filename) bind(C, name="iu_b200_write_vtk") result(rc)
       import
       type(c_ptr), value :: grid
       character(kind=c_char), intent(in) :: filename(*)
       integer(c_int) :: rc
     end function c_write_vtk

     function c_grid_info(grid, info) bind(C, name="iu_b200_grid_info") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int32_t), intent(out) :: info(8)
       integer(c_int) :: rc
     end function c_grid_info

     function c_grid_bounds(grid, rmin, rmax) bind(C, name="iu_b200_grid_bounds") result(rc)
       import
       type(c_ptr), value :: grid
       real(c_double), intent(out) :: rmin(3), rmax(3)
       integer(c_int) :: rc
     end function c_grid_bounds

     function c_get_point_data_index(grid, name) bind(C, name="iu_b200_get_point_data_index") result(ix)
       import
       type(c_ptr), value :: grid
       character(kind=c_char), intent(in) :: name(*)
       integer(c_int) :: ix
     end function c_get_point_data_index

     function c_get_cell_data_index(grid, name) bind(C, name="iu_b200_get_cell_data_index") result(ix)
       import
       type(c_ptr), value :: grid
       character(kind=c_char), intent(in) :: name(*)
       integer(c_int) :: ix
     end function c_get_cell_data_index

     function c_get_icell_data_index(grid, name) bind(C, name="iu_b200_get_icell_data_index") result(ix)
       import
       type(c_ptr), value :: grid
       character(kind=c_char), intent(in) :: name(*)
       integer(c_int) :: ix
     end function c_get_icell_data_index

     function c_add_point_data(grid, name, i_var) bind(C, name="iu_b200_add_point_data") result(rc)
       import
       type(c_ptr), value :: grid
       character(kind=c_char), intent(in) :: name(*)
       integer(c_int), intent(out) :: i_var
       integer(c_int) :: rc
     end function c_add_point_data

     function c_add_cell_data(grid, name, i_var) bind(C, name="iu_b200_add_cell_data") result(rc)
       import
       type(c_ptr), value :: grid
       character(kind=c_char), intent(in) :: name(*)
       integer(c_int), intent(out) :: i_var
       integer(c_int) :: rc
     end function c_add_cell_data

     function c_add_icell_data(grid, name, i_var) bind(C, name="iu_b200_add_icell_data") result(rc)
       import
       type(c_ptr), value :: grid
       character(kind=c_char), intent(in) :: name(*)
       integer(c_int), intent(out) :: i_var
       integer(c_int) :: rc
     end function c_add_icell_data

     function c_update_point_data(grid, col0, n_cols, data) bind(C, name="iu_b200_update_point_data") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int), value :: col0, n_cols
       real(c_double), intent(in) :: data(*)
       integer(c_int) :: rc
     end function c_update_point_data

     function c_update_cell_data(grid, col0, n_cols, data) bind(C, name="iu_b200_update_cell_data") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int), value :: col0, n_cols
       real(c_double), intent(in) :: data(*)
       integer(c_int) :: rc
     end function c_update_cell_data

     function c_update_icell_data(grid, col0, n_cols, data) bind(C, name="iu_b200_update_icell_data") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int), value :: col0, n_cols
       integer(c_int32_t), intent(in) :: data(*)
       integer(c_int) :: rc
     end function c_update_icell_data

     function c_download_point_data(grid, col0, n_cols, data) bind(C, name="iu_b200_download_point_data") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int), value :: col0, n_cols
       real(c_double), intent(out) :: data(*)
       integer(c_int) :: rc
     end function c_download_point_data

     function c_download_cell_data(grid, col0, n_cols, data) bind(C, name="iu_b200_download_cell_data") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int), value :: col0, n_cols
       real(c_double), intent(out) :: data(*)
       integer(c_int) :: rc
     end function c_download_cell_data

     function c_download_icell_data(grid, col0, n_cols, data) bind(C, name="iu_b200_download_icell_data") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int), value :: col0, n_cols
       integer(c_int32_t), intent(out) :: data(*)
       integer(c_int) :: rc
     end function c_download_icell_data

     function c_get_cell_batch(grid, n, r, i_cell) bind(C, name="iu_b200_get_cell_batch") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int64_t), value :: n
       real(c_double), intent(in) :: r(3, *)
       integer(c_int32_t), intent(inout) :: i_cell(*)
       integer(c_int) :: rc
     end function c_get_cell_batch

     function c_interpolate_at_batch(grid, n, r, n_vars, i_vars, var_at_r, i_cell) &
          bind(C, name="iu_b200_interpolate_at_batch") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int64_t), value :: n
       real(c_double), intent(in) :: r(3, *)
       integer(c_int), value :: n_vars
       integer(c_int32_t), intent(in) :: i_vars(*)
       real(c_double), intent(inout) :: var_at_r(n_vars, *)
       integer(c_int32_t), intent(inout) :: i_cell(*)
       integer(c_int) :: rc
     end function c_interpolate_at_batch

     function c_interpolate_at_icell_batch(grid, n, r, n_vars, i_vars, var_at_r, i_cell) &
          bind(C, name="iu_b200_interpolate_at_icell_batch") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int64_t), value :: n
       real(c_double), intent(in) :: r(3, *)
       integer(c_int), value :: n_vars
       integer(c_int32_t), intent(in) :: i_vars(*)
       real(c_double), intent(inout) :: var_at_r(n_vars, *)
       integer(c_int32_t), intent(in) :: i_cell(*)
       integer(c_int) :: rc
     end function c_interpolate_at_icell_batch

     function c_interp_and_cell_scalar_batch(grid, n, r, n_vars, i_vars, i_cell_var, var_at_r, cell_res, i_cell) &
          bind(C, name="iu_b200_interpolate_and_cell_scalar_at_batch") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int64_t), value :: n
       real(c_double), intent(in) :: r(3, *)
       integer(c_int), value :: n_vars
       integer(c_int32_t), intent(in) :: i_vars(*)
       integer(c_int), value :: i_cell_var
       real(c_double), intent(inout) :: var_at_r(n_vars, *)
       real(c_double), intent(inout) :: cell_res(*)
       integer(c_int32_t), intent(inout) :: i_cell(*)
       integer(c_int) :: rc
     end function c_interp_and_cell_scalar_batch

     function c_get_cell_scalar_at_batch(grid, n, r, i_var, i_guess, res) &
          bind(C, name="iu_b200_get_cell_scalar_at_batch") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int64_t), value :: n
       real(c_double), intent(in) :: r(3, *)
       integer(c_int), value :: i_var
       integer(c_int32_t), intent(inout) :: i_guess(*)
       real(c_double), intent(inout) :: res(*)
       integer(c_int) :: rc
     end function c_get_cell_scalar_at_batch

     function c_get_icell_scalar_at_batch(grid, n, r, i_var, i_guess, res) &
          bind(C, name="iu_b200_get_icell_scalar_at_batch") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int64_t), value :: n
       real(c_double), intent(in) :: r(3, *)
       integer(c_int), value :: i_var
       integer(c_int32_t), intent(inout) :: i_guess(*)
       integer(c_int32_t), intent(inout) :: res(*)
       integer(c_int) :: rc
     end function c_get_icell_scalar_at_batch

     function c_walk_batch(grid, n, max_steps, r0, r1, ic0, ic1, r_p, n_steps, status, i_icell_mask) &
          bind(C, name="iu_b200_get_cell_through_neighbors_batch") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int64_t), value :: n
       integer(c_int), value :: max_steps
       real(c_double), intent(in) :: r0(3, *), r1(3, *)
       integer(c_int32_t), intent(in) :: ic0(*)
       integer(c_int32_t), intent(inout) :: ic1(*)
       real(c_double), intent(inout) :: r_p(3, *)
       integer(c_int32_t), intent(inout) :: n_steps(*), status(*)
       integer(c_int), value :: i_icell_mask
       integer(c_int) :: rc
     end function c_walk_batch

     function c_trace_field_batch(grid, n, ndim, nvar, integrand, i_field, min_dx, max_dx, max_steps, rtol, atol, &
          reverse, y, y_field, n_steps, axisymmetric, i_icell_mask, mask_value, boundary_material) &
          bind(C, name="iu_b200_integrate_along_field_batch") result(rc)
       import
       type(c_ptr), value :: grid
       integer(c_int64_t), value :: n
       integer(c_int), value :: ndim, nvar, integrand
       integer(c_int32_t), intent(in) :: i_field(*)
       real(c_double), value :: min_dx, max_dx
       integer(c_int), value :: max_steps
       real(c_double), value :: rtol, atol
       integer(c_int), value :: reverse
       real(c_double), intent(inout) :: y(ndim + nvar, max_steps, *)
       real(c_double), intent(inout) :: y_field(ndim, max_steps, *)
       integer(c_int32_t), intent(inout) :: n_steps(*)
       integer(c_int), value :: axisymmetric, i_icell_mask, mask_value
       integer(c_int32_t), intent(inout) :: boundary_material(*)
       integer(c_int) :: rc
     end function c_trace_field_batch
  end interface

  ! Public types
  public :: iu_grid_t
  public :: iu_triangle, iu_quad, iu_tetra
  public :: iu_ndim_cell_type

  ! Public methods with the reference's names and argument lists
  public :: iu_read_grid
  public :: iu_get_point_data_index
  public :: iu_get_cell_data_index
  public :: iu_get_icell_data_index
  public :: iu_add_cell_data
  public :: iu_add_icell_data
  public :: iu_add_point_data
  public :: iu_get_cell
  public :: iu_get_cell_through_neighbors
  public :: iu_get_cell_scalar_at
  public :: iu_get_icell_scalar_at
  public :: iu_interpolate_at
  public :: iu_interpolate_scalar_at

  ! New: batched variants, device refresh, lifecycle
  public :: iu_get_cell_batch
  public :: iu_get_cell_through_neighbors_batch
  public :: iu_get_cell_scalar_at_batch
  public :: iu_interpolate_and_cell_scalar_at_batch
  public :: iu_get_icell_scalar_at_batch
  public :: iu_interpolate_at_batch
  public :: iu_interpolate_scalar_at_batch
  public :: iu_interpolate_at_icell_batch
  public :: iu_integrate_along_field_batch
  public :: iu_sync_point_data
  public :: iu_sync_cell_data
  public :: iu_sync_icell_data
  public :: iu_destroy_grid
  public :: iu_device_count
  public :: iu_get_cell_centers
  public :: iu_write_vtk

contains

  !> Abort with the library's message (the reference uses `error stop` for the same conditions)
  subroutine check(rc, where)
    integer(c_int), intent(in)   :: rc
    character(len=*), intent(in) :: where
    character(kind=c_char), pointer :: msg(:)
    type(c_ptr) :: p
    integer :: n

    if (rc == 0) return
    p = c_last_error()
    if (c_associated(p)) then
       call c_f_pointer(p, msg, [1024])
       n = 0
       do while (n < 1024)
          if (msg(n+1) == c_null_char) exit
          n = n + 1
       end do
       write(error_unit, *) where, ": ", msg(1:n)
    end if
    error stop "iu_b200 call failed"
  end subroutine check

  pure function c_string(s) result(c)
    character(len=*), intent(in) :: s
    character(kind=c_char)       :: c(len_trim(s)+1)
    integer :: i
    do i = 1, len_trim(s)
       c(i) = s(i:i)
    end do
    c(len_trim(s)+1) = c_null_char
  end function c_string

  !> iu_read_grid, m_interp_unstructured.f90:820-927 (*.binda files only: the vtu -> binda
  !> conversion of :788-818 is a separate tool)
  subroutine iu_read_grid(filename, ug, coord_scale_factor, device, devices)
    character(len=*), intent(in)   :: filename
    type(iu_grid_t), intent(inout) :: ug
    real(dp), intent(in), optional :: coord_scale_factor
    integer, intent(in), optional  :: device !< CUDA device ordinal (default: current)
    !> several GPUs: devices(1) builds the mesh, the others receive a device-to-device replica;
    !> the *_batch calls then split their queries into one slab per GPU
    integer, intent(in), optional  :: devices(:)
    real(c_double) :: scale
    integer(c_int) :: dev
    integer(c_int32_t) :: info(8)

    scale = 1.0_dp; if (present(coord_scale_factor)) scale = coord_scale_factor
    dev = -1; if (present(device)) dev = device
    if (present(devices)) then
       if (size(devices) > 0) dev = devices(1)
    end if
    if (c_associated(ug%handle)) call iu_destroy_grid(ug)
    call check(c_grid_from_binda(c_string(filename), scale, dev, ug%handle), "iu_read_grid")
    if (present(devices)) then
       if (size(devices) > 1) call check(c_grid_replicate(ug%handle, size(devices), devices), "iu_read_grid")
    end if
    call check(c_grid_info(ug%handle, info), "iu_read_grid")
    ug%cell_type = info(1)
    ug%n_points = info(2)
    ug%n_cells = info(3)
    ug%n_points_per_cell = info(4)
    ug%n_faces_per_cell = info(4)
    ug%n_point_data = info(5)
    ug%n_cell_data = info(6)
    ug%n_icell_data = info(7)
    call check(c_grid_bounds(ug%handle, ug%rmin, ug%rmax), "iu_read_grid")

    ! host mirrors of the data tables, as the reference exposes them
    if (allocated(ug%point_data)) deallocate(ug%point_data)
    if (allocated(ug%cell_data)) deallocate(ug%cell_data)
    if (allocated(ug%icell_data)) deallocate(ug%icell_data)
    allocate(ug%point_data(ug%n_points, ug%n_point_data))
    allocate(ug%cell_data(ug%n_cells, ug%n_cell_data))
    allocate(ug%icell_data(ug%n_cells, ug%n_icell_data))
    if (ug%n_point_data > 0) call check(c_download_point_data(ug%handle, 1, &
         ug%n_point_data, ug%point_data), "iu_read_grid")
    if (ug%n_cell_data > 0) call check(c_download_cell_data(ug%handle, 1, &
         ug%n_cell_data, ug%cell_data), "iu_read_grid")
    if (ug%n_icell_data > 0) call check(c_download_icell_data(ug%handle, 1, &
         ug%n_icell_data, ug%icell_data), "iu_read_grid")
  end subroutine iu_read_grid

  !> number of GPUs that hold a replica of the mesh
  integer function iu_device_count(ug)
    type(iu_grid_t), intent(in) :: ug
    iu_device_count = c_device_count(ug%handle)
  end function iu_device_count

  !> iu_get_cell_center (:264-269) for every cell: centers(3, n_cells), the values the search uses
  subroutine iu_get_cell_centers(ug, centers)
    type(iu_grid_t), intent(in) :: ug
    real(dp), intent(out)       :: centers(3, ug%n_cells)
    call check(c_export_geometry(ug%handle, c_null_ptr, c_null_ptr, c_null_ptr, centers), "iu_get_cell_centers")
  end subroutine iu_get_cell_centers

  subroutine iu_destroy_grid(ug)
    type(iu_grid_t), intent(inout) :: ug
    integer(c_int) :: rc
    rc = c_grid_destroy(ug%handle)
    ug%handle = c_null_ptr
  end subroutine iu_destroy_grid

  !> iu_write_vtk (:930-985): the device-resident grid with its current data as a .vtu file
  !> (call iu_sync_*_data first if the host copies were modified)
  subroutine iu_write_vtk(ug, filename)
    type(iu_grid_t), intent(in)  :: ug
    character(len=*), intent(in) :: filename
    call check(c_write_vtk(ug%handle, c_string(trim(filename))), "iu_write_vtk")
  end subroutine iu_write_vtk

  !> :106-116
  subroutine iu_get_point_data_index(ug, name, ix)
    type(iu_grid_t), intent(in)  :: ug
    character(len=*), intent(in) :: name
    integer, intent(out)         :: ix
    ix = c_get_point_data_index(ug%handle, c_string(name))
  end subroutine iu_get_point_data_index

  !> :224-234
  subroutine iu_get_cell_data_index(ug, name, ix)
    type(iu_grid_t), intent(in)  :: ug
    character(len=*), intent(in) :: name
    integer, intent(out)         :: ix
    ix = c_get_cell_data_index(ug%handle, c_string(name))
  end subroutine iu_get_cell_data_index

  !> :237-247
  subroutine iu_get_icell_data_index(ug, name, ix)
    type(iu_grid_t), intent(in)  :: ug
    character(len=*), intent(in) :: name
    integer, intent(out)         :: ix
    ix = c_get_icell_data_index(ug%handle, c_string(name))
  end subroutine iu_get_icell_data_index

  !> :149-161: append a zero column; the host mirror grows with it
  subroutine iu_add_point_data(ug, name, i_var)
    type(iu_grid_t), intent(inout) :: ug
    character(len=*), intent(in)   :: name
    integer, intent(out)           :: i_var
    real(dp), allocatable :: tmp(:, :)
    integer(c_int) :: iv

    call check(c_add_point_data(ug%handle, c_string(name), iv), "iu_add_point_data")
    i_var = iv
    allocate(tmp(ug%n_points, i_var))
    tmp(:, 1:ug%n_point_data) = ug%point_data(:, 1:ug%n_point_data)
    tmp(:, i_var) = 0.0_dp
    call move_alloc(tmp, ug%point_data)
    ug%n_point_data = i_var
  end subroutine iu_add_point_data

  !> :119-131
  subroutine iu_add_cell_data(ug, name, i_var)
    type(iu_grid_t), intent(inout) :: ug
    character(len=*), intent(in)   :: name
    integer, intent(out)           :: i_var
    real(dp), allocatable :: tmp(:, :)
    integer(c_int) :: iv

    call check(c_add_cell_data(ug%handle, c_string(name), iv), "iu_add_cell_data")
    i_var = iv
    allocate(tmp(ug%n_cells, i_var))
    tmp(:, 1:ug%n_cell_data) = ug%cell_data(:, 1:ug%n_cell_data)
    tmp(:, i_var) = 0.0_dp
    call move_alloc(tmp, ug%cell_data)
    ug%n_cell_data = i_var
  end subroutine iu_add_cell_data

  !> :134-146
  subroutine iu_add_icell_data(ug, name, i_var)
    type(iu_grid_t), intent(inout) :: ug
    character(len=*), intent(in)   :: name
    integer, intent(out)           :: i_var
    integer, allocatable :: tmp(:, :)
    integer(c_int) :: iv

    call check(c_add_icell_data(ug%handle, c_string(name), iv), "iu_add_icell_data")
    i_var = iv
    allocate(tmp(ug%n_cells, i_var))
    tmp(:, 1:ug%n_icell_data) = ug%icell_data(:, 1:ug%n_icell_data)
    tmp(:, i_var) = 0
    call move_alloc(tmp, ug%icell_data)
    ug%n_icell_data = i_var
  end subroutine iu_add_icell_data

  !> Push ug%point_data (all columns, or i_var only) to the device after the caller changed it
  subroutine iu_sync_point_data(ug, i_var)
    type(iu_grid_t), intent(in)   :: ug
    integer, intent(in), optional :: i_var
    if (present(i_var)) then
       call check(c_update_point_data(ug%handle, i_var, 1, ug%point_data(:, i_var)), "iu_sync_point_data")
    else if (ug%n_point_data > 0) then
       call check(c_update_point_data(ug%handle, 1, ug%n_point_data, ug%point_data), "iu_sync_point_data")
    end if
  end subroutine iu_sync_point_data

  subroutine iu_sync_cell_data(ug, i_var)
    type(iu_grid_t), intent(in)   :: ug
    integer, intent(in), optional :: i_var
    if (present(i_var)) then
       call check(c_update_cell_data(ug%handle, i_var, 1, ug%cell_data(:, i_var)), "iu_sync_cell_data")
    else if (ug%n_cell_data > 0) then
       call check(c_update_cell_data(ug%handle, 1, ug%n_cell_data, ug%cell_data), "iu_sync_cell_data")
    end if
  end subroutine iu_sync_cell_data

  subroutine iu_sync_icell_data(ug, i_var)
    type(iu_grid_t), intent(in)   :: ug
    integer, intent(in), optional :: i_var
    if (present(i_var)) then
       call check(c_update_icell_data(ug%handle, i_var, 1, ug%icell_data(:, i_var)), "iu_sync_icell_data")
    else if (ug%n_icell_data > 0) then
       call check(c_update_icell_data(ug%handle, 1, ug%n_icell_data, ug%icell_data), "iu_sync_icell_data")
    end if
  end subroutine iu_sync_icell_data

  ! ---- batched variants -------------------------------------------------------------------

  !> iu_get_cell (:414-434) for n points
  subroutine iu_get_cell_batch(ug, n, r, i_cell)
    type(iu_grid_t), intent(in)    :: ug
    integer(int64), intent(in)     :: n
    real(dp), intent(in)           :: r(3, n)
    integer, intent(inout)         :: i_cell(n)
    call check(c_get_cell_batch(ug%handle, n, r, i_cell), "iu_get_cell")
  end subroutine iu_get_cell_batch

  !> iu_interpolate_at (:480-495) for n points
  subroutine iu_interpolate_at_batch(ug, n, r, n_vars, i_vars, var_at_r, i_cell)
    type(iu_grid_t), intent(in) :: ug
    integer(int64), intent(in)  :: n
    real(dp), intent(in)        :: r(3, n)
    integer, intent(in)         :: n_vars
    integer, intent(in)         :: i_vars(n_vars)
    real(dp), intent(inout)     :: var_at_r(n_vars, n)
    integer, intent(inout)      :: i_cell(n)
    call check(c_interpolate_at_batch(ug%handle, n, r, n_vars, i_vars, var_at_r, i_cell), "iu_interpolate_at")
  end subroutine iu_interpolate_at_batch

  !> iu_interpolate_scalar_at (:465-477) for n points
  subroutine iu_interpolate_scalar_at_batch(ug, n, r, i_var, var_at_r, i_cell)
    type(iu_grid_t), intent(in) :: ug
    integer(int64), intent(in)  :: n
    real(dp), intent(in)        :: r(3, n)
    integer, intent(in)         :: i_var
    real(dp), intent(inout)     :: var_at_r(n)
    integer, intent(inout)      :: i_cell(n)
    integer(c_int32_t) :: iv(1)
    iv(1) = i_var
    call check(c_interpolate_at_batch(ug%handle, n, r, 1, iv, var_at_r, i_cell), "iu_interpolate_scalar_at")
  end subroutine iu_interpolate_scalar_at_batch

  !> iu_interpolate_at_icell (:498-527) for n points whose cells are known
  subroutine iu_interpolate_at_icell_batch(ug, n, r, n_vars, i_vars, var_at_r, i_cell)
    type(iu_grid_t), intent(in) :: ug
    integer(int64), intent(in)  :: n
    real(dp), intent(in)        :: r(3, n)
    integer, intent(in)         :: n_vars
    integer, intent(in)         :: i_vars(n_vars)
    real(dp), intent(inout)     :: var_at_r(n_vars, n)
    integer, intent(in)         :: i_cell(n)
    call check(c_interpolate_at_icell_batch(ug%handle, n, r, n_vars, i_vars, var_at_r, i_cell), &
         "iu_interpolate_at_icell")
  end subroutine iu_interpolate_at_icell_batch

  !> iu_interpolate_at (:480-495) followed by iu_get_cell_scalar_at (:437-448) with the cell just
  !> found as i_guess, for n points in one pass over the queries
  subroutine iu_interpolate_and_cell_scalar_at_batch(ug, n, r, n_vars, i_vars, i_cell_var, var_at_r, cell_res, i_cell)
    type(iu_grid_t), intent(in) :: ug
    integer(int64), intent(in)  :: n
    real(dp), intent(in)        :: r(3, n)
    integer, intent(in)         :: n_vars
    integer, intent(in)         :: i_vars(n_vars)
    integer, intent(in)         :: i_cell_var
    real(dp), intent(inout)     :: var_at_r(n_vars, n)
    real(dp), intent(inout)     :: cell_res(n)
    integer, intent(inout)      :: i_cell(n)
    call check(c_interp_and_cell_scalar_batch(ug%handle, n, r, n_vars, i_vars, i_cell_var, var_at_r, cell_res, i_cell), &
         "iu_interpolate_and_cell_scalar_at")
  end subroutine iu_interpolate_and_cell_scalar_at_batch

  !> iu_get_cell_scalar_at (:437-448) for n points
  subroutine iu_get_cell_scalar_at_batch(ug, n, r, i_var, i_guess, res)
    type(iu_grid_t), intent(in) :: ug
    integer(int64), intent(in)  :: n
    real(dp), intent(in)        :: r(3, n)
    integer, intent(in)         :: i_var
    integer, intent(inout)      :: i_guess(n)
    real(dp), intent(inout)     :: res(n)
    call check(c_get_cell_scalar_at_batch(ug%handle, n, r, i_var, i_guess, res), "iu_get_cell_scalar_at")
  end subroutine iu_get_cell_scalar_at_batch

  !> iu_get_icell_scalar_at (:451-462) for n points
  subroutine iu_get_icell_scalar_at_batch(ug, n, r, i_var, i_guess, res)
    type(iu_grid_t), intent(in) :: ug
    integer(int64), intent(in)  :: n
    real(dp), intent(in)        :: r(3, n)
    integer, intent(in)         :: i_var
    integer, intent(inout)      :: i_guess(n)
    integer, intent(inout)      :: res(n)
    call check(c_get_icell_scalar_at_batch(ug%handle, n, r, i_var, i_guess, res), "iu_get_icell_scalar_at")
  end subroutine iu_get_icell_scalar_at_batch

  !> iu_get_cell_through_neighbors (:667-725) for n rays
  subroutine iu_get_cell_through_neighbors_batch(ug, n, max_steps, r0, r1, ic0, ic1, r_p, &
       n_steps, status, i_icell_mask)
    type(iu_grid_t), intent(in)   :: ug
    integer(int64), intent(in)    :: n
    integer, intent(in)           :: max_steps
    real(dp), intent(in)          :: r0(3, n), r1(3, n)
    integer, intent(in)           :: ic0(n)
    integer, intent(inout)        :: ic1(n)
    real(dp), intent(inout)       :: r_p(3, n)
    integer, intent(inout)        :: n_steps(n), status(n)
    integer, intent(in), optional :: i_icell_mask
    integer(c_int) :: mask, cap

    mask = 0; if (present(i_icell_mask)) mask = i_icell_mask
    cap = min(max_steps, huge(1) - 1)   ! the reference passes huge(1), :431
    call check(c_walk_batch(ug%handle, n, cap, r0, r1, ic0, ic1, r_p, n_steps, status, mask), &
         "iu_get_cell_through_neighbors")
  end subroutine iu_get_cell_through_neighbors_batch

  !> iu_integrate_along_field (:989-1217) for n field lines.  Same argument order as the reference
  !> except that the callback sub_int is replaced by `integrand` (0 none with nvar = 0, 1 dy = 1 as
  !> in test_trace_field.f90:57-64, 2 dy = norm2(field)) and every per-line argument gains a
  !> trailing dimension n.
  subroutine iu_integrate_along_field_batch(ug, n, ndim, nvar, integrand, i_field, &
       min_dx, max_dx, max_steps, rtol, atol, reverse, y, y_field, &
       n_steps, axisymmetric, i_icell_mask, mask_value, boundary_material)
    type(iu_grid_t), intent(in)   :: ug
    integer(int64), intent(in)    :: n
    integer, intent(in)           :: ndim, nvar, integrand
    integer, intent(in)           :: i_field(ndim)
    real(dp), intent(in)          :: min_dx, max_dx
    integer, intent(in)           :: max_steps
    real(dp), intent(in)          :: rtol, atol
    logical, intent(in)           :: reverse
    real(dp), intent(inout)       :: y(ndim+nvar, max_steps, n)
    real(dp), intent(inout)       :: y_field(ndim, max_steps, n)
    integer, intent(out)          :: n_steps(n)
    logical, intent(in)           :: axisymmetric
    integer, intent(in), optional :: i_icell_mask
    integer, intent(in), optional :: mask_value
    integer, intent(inout), optional :: boundary_material(n)
    integer(c_int) :: mask, mval, axi, rev
    integer, allocatable :: bm(:)

    if (present(i_icell_mask) .neqv. present(mask_value)) &
         error stop "present(i_icell_mask) .neqv. present(mask_value)"     ! :1038-1039
    mask = 0; mval = 0
    if (present(i_icell_mask)) then
       mask = i_icell_mask; mval = mask_value
    end if
    axi = 0; if (axisymmetric) axi = 1
    rev = 0; if (reverse) rev = 1
    if (present(boundary_material)) then
       call check(c_trace_field_batch(ug%handle, n, ndim, nvar, integrand, i_field, min_dx, max_dx, &
            max_steps, rtol, atol, rev, y, y_field, n_steps, axi, mask, mval, boundary_material), &
            "iu_integrate_along_field")
    else
       allocate(bm(n))
       call check(c_trace_field_batch(ug%handle, n, ndim, nvar, integrand, i_field, min_dx, max_dx, &
            max_steps, rtol, atol, rev, y, y_field, n_steps, axi, mask, mval, bm), &
            "iu_integrate_along_field")
    end if
  end subroutine iu_integrate_along_field_batch

  ! ---- the reference's scalar entry points: batches of one --------------------------------

  !> :414-434
  subroutine iu_get_cell(ug, r, i_cell)
    type(iu_grid_t), intent(in) :: ug
    real(dp), intent(in)        :: r(3)
    integer, intent(inout)      :: i_cell
    real(dp) :: r1(3, 1)
    integer  :: ic(1)
    r1(:, 1) = r; ic(1) = i_cell
    call iu_get_cell_batch(ug, 1_int64, r1, ic)
    i_cell = ic(1)
  end subroutine iu_get_cell

  !> :480-495
  subroutine iu_interpolate_at(ug, r, n_vars, i_vars, var_at_r, i_cell)
    type(iu_grid_t), intent(in) :: ug
    real(dp), intent(in)        :: r(3)
    integer, intent(in)         :: n_vars
    integer, intent(in)         :: i_vars(n_vars)
    real(dp), intent(inout)     :: var_at_r(n_vars)
    integer, intent(inout)      :: i_cell
    real(dp) :: r1(3, 1), v(n_vars, 1)
    integer  :: ic(1)
    r1(:, 1) = r; ic(1) = i_cell; v(:, 1) = var_at_r
    call iu_interpolate_at_batch(ug, 1_int64, r1, n_vars, i_vars, v, ic)
    i_cell = ic(1); var_at_r = v(:, 1)
  end subroutine iu_interpolate_at

  !> :465-477
  subroutine iu_interpolate_scalar_at(ug, r, i_var, var_at_r, i_cell)
    type(iu_grid_t), intent(in) :: ug
    real(dp), intent(in)        :: r(3)
    integer, intent(in)         :: i_var
    real(dp), intent(out)       :: var_at_r
    integer, intent(inout)      :: i_cell
    real(dp) :: r1(3, 1), v(1)
    integer  :: ic(1)
    r1(:, 1) = r; ic(1) = i_cell; v(1) = 0.0_dp
    call iu_interpolate_scalar_at_batch(ug, 1_int64, r1, i_var, v, ic)
    i_cell = ic(1); var_at_r = v(1)
  end subroutine iu_interpolate_scalar_at

  !> :437-448
  subroutine iu_get_cell_scalar_at(ug, r, i_var, i_guess, res)
    type(iu_grid_t), intent(in) :: ug
    real(dp), intent(in)        :: r(3)
    integer, intent(in)         :: i_var
    integer, intent(inout)      :: i_guess
    real(dp), intent(inout)     :: res
    real(dp) :: r1(3, 1), v(1)
    integer  :: ic(1)
    r1(:, 1) = r; ic(1) = i_guess; v(1) = res
    call iu_get_cell_scalar_at_batch(ug, 1_int64, r1, i_var, ic, v)
    i_guess = ic(1); res = v(1)
  end subroutine iu_get_cell_scalar_at

  !> :451-462
  subroutine iu_get_icell_scalar_at(ug, r, i_var, i_guess, res)
    type(iu_grid_t), intent(in) :: ug
    real(dp), intent(in)        :: r(3)
    integer, intent(in)         :: i_var
    integer, intent(inout)      :: i_guess
    integer, intent(inout)      :: res
    real(dp) :: r1(3, 1)
    integer  :: ic(1), v(1)
    r1(:, 1) = r; ic(1) = i_guess; v(1) = res
    call iu_get_icell_scalar_at_batch(ug, 1_int64, r1, i_var, ic, v)
    i_guess = ic(1); res = v(1)
  end subroutine iu_get_icell_scalar_at

  !> :667-725
  subroutine iu_get_cell_through_neighbors(ug, max_steps, r0, r1, ic0, ic1, r_p, n_steps, status, i_icell_mask)
    type(iu_grid_t), intent(in)   :: ug
    integer, intent(in)           :: max_steps
    real(dp), intent(in)          :: r0(3), r1(3)
    integer, intent(in)           :: ic0
    integer, intent(out)          :: ic1
    real(dp), intent(out)         :: r_p(3)
    integer, intent(out)          :: n_steps
    integer, intent(out)          :: status
    integer, intent(in), optional :: i_icell_mask
    real(dp) :: a(3, 1), b(3, 1), p(3, 1)
    integer  :: c0(1), c1(1), ns(1), st(1), mask

    mask = 0; if (present(i_icell_mask)) mask = i_icell_mask
    a(:, 1) = r0; b(:, 1) = r1; c0(1) = ic0
    call iu_get_cell_through_neighbors_batch(ug, 1_int64, max_steps, a, b, c0, c1, p, ns, st, mask)
    ic1 = c1(1); r_p = p(:, 1); n_steps = ns(1); status = st(1)
  end subroutine iu_get_cell_through_neighbors

end module m_iu_b200

!> The reference's module name as a thin re-export: user code keeps its `use m_interp_unstructured`
!> line (m_interp_unstructured.f90:1) and links libiu_b200.so instead of libinterp_unstructured.a.
!> Public procedures of the reference that are intentionally absent: iu_point_is_inside_cell,
!> iu_reserve_*_storage, and the callback form of iu_integrate_along_field (INTEGRATION.md, section 3).
module m_interp_unstructured
  use m_iu_b200
  implicit none
  public
end module m_interp_unstructured
