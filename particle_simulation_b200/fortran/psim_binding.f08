! psim_binding.f08 — ISO_C_BINDING interface to libpsim_b200.so (include/psim.h).
!
! UNCOMPILED ON THIS PROJECT'S MACHINES: neither the build container nor the GPU
! box has a Fortran compiler (SURVEY.md F3).  The C++ host (csrc/host/) and the
! Python host (native.py) call the same entry points and are tested; this file
! is what a maintainer of the reference adds to
! fortran_src/program_particle_simulation_openmp.f08 to use the B200 path.
!
! Usage sketch (replaces program_particle_simulation_openmp.f08:59-123):
!
!   type(psim_params) :: p
!   type(c_ptr) :: sim
!   real(rt), target :: pos(3, n), vel(3, n)       ! packed xyz, particle-major
!   integer(c_int32_t), target :: id(n)
!   integer(c_int8_t), target :: active(n)
!   p = psim_params(delta_t, epsilon, sigma, mass, r_cut, length_of_area, length_of_cell, nb_of_threads, 0)
!   ierr = psim_create(p, int(storage_size(1.0_rt) / 8, c_int), int(n, c_int64_t), &
!                      c_loc(id), c_loc(pos), c_loc(vel), -1_c_int, sim)
!   do frame = 1, max_time_step / write_time_step
!     ierr = psim_advance(sim, int(write_time_step, c_int64_t))
!     ierr = psim_get_frame(sim, c_loc(id), c_loc(pos), c_loc(vel), c_loc(active))
!     call write_data(...)                         ! as before, skipping active(i) == 0
!   end do
!   ierr = psim_destroy(sim)
module psim_binding
  use, intrinsic :: iso_c_binding
  implicit none
  private
  public :: psim_params, psim_create, psim_destroy, psim_advance, psim_synchronize, psim_get_frame, &
            psim_get_aux, psim_set_option, psim_set_positions, psim_status_string_ptr

  integer(c_int), parameter, public :: PSIM_OK = 0
  integer(c_int), parameter, public :: PSIM_OPT_FORCE_ALGO = 0, PSIM_FORCE_ALL_PAIRS = 0, PSIM_FORCE_CELL_BINNED = 1

  type, bind(C) :: psim_params
    real(c_double) :: delta_t, epsilon, sigma, mass, r_cut_factor
    real(c_double) :: length_of_area(3), length_of_cell(3)
    integer(c_int32_t) :: nb_of_threads, reserved
  end type psim_params

  interface
    integer(c_int) function psim_create(params, precision, n, id, pos, vel, device, sim) bind(C, name="psim_create")
      import :: c_int, c_int64_t, c_ptr, psim_params
      type(psim_params), intent(in) :: params
      integer(c_int), value :: precision
      integer(c_int64_t), value :: n
      type(c_ptr), value :: id, pos, vel
      integer(c_int), value :: device
      type(c_ptr), intent(out) :: sim
    end function psim_create

    integer(c_int) function psim_destroy(sim) bind(C, name="psim_destroy")
      import :: c_int, c_ptr
      type(c_ptr), value :: sim
    end function psim_destroy

    integer(c_int) function psim_set_option(sim, option, value) bind(C, name="psim_set_option")
      import :: c_int, c_int64_t, c_ptr
      type(c_ptr), value :: sim
      integer(c_int), value :: option
      integer(c_int64_t), value :: value
    end function psim_set_option

    integer(c_int) function psim_advance(sim, nsteps) bind(C, name="psim_advance")
      import :: c_int, c_int64_t, c_ptr
      type(c_ptr), value :: sim
      integer(c_int64_t), value :: nsteps
    end function psim_advance

    integer(c_int) function psim_synchronize(sim) bind(C, name="psim_synchronize")
      import :: c_int, c_ptr
      type(c_ptr), value :: sim
    end function psim_synchronize

    integer(c_int) function psim_get_frame(sim, id, pos, vel, active) bind(C, name="psim_get_frame")
      import :: c_int, c_ptr
      type(c_ptr), value :: sim, id, pos, vel, active
    end function psim_get_frame

    integer(c_int) function psim_get_aux(sim, displacement, force_old, force_new) bind(C, name="psim_get_aux")
      import :: c_int, c_ptr
      type(c_ptr), value :: sim, displacement, force_old, force_new
    end function psim_get_aux

    integer(c_int) function psim_set_positions(sim, pos, active) bind(C, name="psim_set_positions")
      import :: c_int, c_ptr
      type(c_ptr), value :: sim, pos, active
    end function psim_set_positions

    type(c_ptr) function psim_status_string_ptr(status) bind(C, name="psim_status_string")
      import :: c_int, c_ptr
      integer(c_int), value :: status
    end function psim_status_string_ptr
  end interface
end module psim_binding
