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
  public :: psim_params, psim_counters, psim_create, psim_create_sharded, psim_comm_unique_id, psim_destroy, &
            psim_advance, psim_advance_timed, psim_synchronize, psim_get_frame, psim_get_aux, psim_get_counters, &
            psim_get_shard, psim_set_option, psim_set_positions, psim_snapshot_async, psim_snapshot_wait, &
            psim_compute_forces_host, psim_count_pairs_in_range, psim_energy, psim_compute_locations_static, &
            psim_measure_fma_peak, psim_measure_tensor_peak, psim_status_string_ptr, psim_last_error_ptr, &
            psim_version_ptr

  ! psim_status
  integer(c_int), parameter, public :: PSIM_OK = 0, PSIM_ERR_INVALID_ARGUMENT = 1, PSIM_ERR_NO_DEVICE = 2, &
                                       PSIM_ERR_CUDA = 3, PSIM_ERR_NO_PARTICLES = 4, PSIM_ERR_UNSUPPORTED = 5, &
                                       PSIM_ERR_COMM = 6, PSIM_ERR_IO = 7
  ! psim_option and its values
  integer(c_int), parameter, public :: PSIM_OPT_FORCE_ALGO = 0, PSIM_OPT_FUSED_STEP = 1, PSIM_OPT_KERNEL_TIMING = 2, &
                                       PSIM_OPT_ALLPAIRS_KERNEL = 3
  integer(c_int64_t), parameter, public :: PSIM_FORCE_ALL_PAIRS = 0, PSIM_FORCE_CELL_BINNED = 1
  integer(c_int64_t), parameter, public :: PSIM_ALLPAIRS_TENSOR = 0, PSIM_ALLPAIRS_DIFFERENCE_FORM = 1, &
                                           PSIM_ALLPAIRS_CUDA_CORE = 5
  integer(c_int), parameter, public :: PSIM_FRAME_SLOTS = 4
  integer(c_int64_t), parameter, public :: PSIM_STATIC_MAX_PARTICLES = 1048576_c_int64_t

  type, bind(C) :: psim_params
    real(c_double) :: delta_t, epsilon, sigma, mass, r_cut_factor
    real(c_double) :: length_of_area(3), length_of_cell(3)
    integer(c_int32_t) :: nb_of_threads, reserved
  end type psim_params

  type, bind(C) :: psim_counters
    integer(c_int64_t) :: steps, rebuilds, kernel_launches, force_launches
    real(c_double)     :: force_ms
    integer(c_int64_t) :: force_timed, n_active, filter_pairs, filter_tests, peer_push_steps
  end type psim_counters

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

    ! One i-slice per process (the reference's thread slices, one per GPU).  nccl_unique_id: the 128 bytes that
    ! psim_comm_unique_id produced on rank 0, distributed by the host (MPI_Bcast, a file ...).
    integer(c_int) function psim_create_sharded(params, precision, n, id, pos, vel, device, rank, world, &
                                                nccl_unique_id, sim) bind(C, name="psim_create_sharded")
      import :: c_int, c_int64_t, c_ptr, psim_params
      type(psim_params), intent(in) :: params
      integer(c_int), value :: precision
      integer(c_int64_t), value :: n
      type(c_ptr), value :: id, pos, vel
      integer(c_int), value :: device, rank, world
      type(c_ptr), value :: nccl_unique_id
      type(c_ptr), intent(out) :: sim
    end function psim_create_sharded

    integer(c_int) function psim_comm_unique_id(out128) bind(C, name="psim_comm_unique_id")
      import :: c_int, c_ptr
      type(c_ptr), value :: out128
    end function psim_comm_unique_id

    integer(c_int) function psim_advance_timed(sim, nsteps, elapsed_ms) bind(C, name="psim_advance_timed")
      import :: c_int, c_int64_t, c_double, c_ptr
      type(c_ptr), value :: sim
      integer(c_int64_t), value :: nsteps
      real(c_double), intent(out) :: elapsed_ms
    end function psim_advance_timed

    integer(c_int) function psim_get_counters(sim, counters) bind(C, name="psim_get_counters")
      import :: c_int, c_ptr, psim_counters
      type(c_ptr), value :: sim
      type(psim_counters), intent(out) :: counters
    end function psim_get_counters

    integer(c_int) function psim_get_shard(sim, i_lo, i_hi) bind(C, name="psim_get_shard")
      import :: c_int, c_int64_t, c_ptr
      type(c_ptr), value :: sim
      integer(c_int64_t), intent(out) :: i_lo, i_hi      ! 0-based, half-open
    end function psim_get_shard

    integer(c_int) function psim_snapshot_async(sim, slot) bind(C, name="psim_snapshot_async")
      import :: c_int, c_ptr
      type(c_ptr), value :: sim
      integer(c_int), value :: slot
    end function psim_snapshot_async

    integer(c_int) function psim_snapshot_wait(sim, slot, id, pos, vel, active) bind(C, name="psim_snapshot_wait")
      import :: c_int, c_ptr
      type(c_ptr), value :: sim
      integer(c_int), value :: slot
      type(c_ptr), intent(out) :: id, pos, vel, active   ! pointers into the slot's pinned buffers: c_f_pointer them
    end function psim_snapshot_wait

    integer(c_int) function psim_compute_forces_host(params, precision, n, pos, active, algo, device, force, &
                                                     pairs_in_range) bind(C, name="psim_compute_forces_host")
      import :: c_int, c_int64_t, c_ptr, psim_params
      type(psim_params), intent(in) :: params
      integer(c_int), value :: precision
      integer(c_int64_t), value :: n
      type(c_ptr), value :: pos, active
      integer(c_int), value :: algo, device
      type(c_ptr), value :: force
      integer(c_int64_t), intent(out) :: pairs_in_range
    end function psim_compute_forces_host

    integer(c_int) function psim_count_pairs_in_range(sim, pairs) bind(C, name="psim_count_pairs_in_range")
      import :: c_int, c_int64_t, c_ptr
      type(c_ptr), value :: sim
      integer(c_int64_t), intent(out) :: pairs
    end function psim_count_pairs_in_range

    integer(c_int) function psim_energy(sim, kinetic, potential, momentum) bind(C, name="psim_energy")
      import :: c_int, c_double, c_ptr
      type(c_ptr), value :: sim
      real(c_double), intent(out) :: kinetic, potential, momentum(3)
    end function psim_energy

    ! The drift step on DEVICE arrays the caller owns (c_devptr / device addresses passed as c_ptr), for hosts
    ! that cannot let a library allocate GPU memory; see include/psim.h for its limits.
    integer(c_int) function psim_compute_locations_static(params, precision, n, pos_dev, vel_dev, force_old_dev, &
                                                          force_new_dev, displacement_dev, active_dev, &
                                                          rebuild_flag_dev, cuda_stream) &
                                                          bind(C, name="psim_compute_locations_static")
      import :: c_int, c_int64_t, c_ptr, psim_params
      type(psim_params), intent(in) :: params
      integer(c_int), value :: precision
      integer(c_int64_t), value :: n
      type(c_ptr), value :: pos_dev, vel_dev, force_old_dev, force_new_dev, displacement_dev, active_dev
      type(c_ptr), value :: rebuild_flag_dev, cuda_stream
    end function psim_compute_locations_static

    integer(c_int) function psim_measure_fma_peak(precision, device, tflops, sm_mhz) bind(C, name="psim_measure_fma_peak")
      import :: c_int, c_double
      integer(c_int), value :: precision, device
      real(c_double), intent(out) :: tflops, sm_mhz
    end function psim_measure_fma_peak

    integer(c_int) function psim_measure_tensor_peak(kind, device, tflops, sm_mhz) bind(C, name="psim_measure_tensor_peak")
      import :: c_int, c_double
      integer(c_int), value :: kind, device
      real(c_double), intent(out) :: tflops, sm_mhz
    end function psim_measure_tensor_peak

    type(c_ptr) function psim_last_error_ptr() bind(C, name="psim_last_error")
      import :: c_ptr
    end function psim_last_error_ptr

    type(c_ptr) function psim_version_ptr() bind(C, name="psim_version")
      import :: c_ptr
    end function psim_version_ptr

    type(c_ptr) function psim_status_string_ptr(status) bind(C, name="psim_status_string")
      import :: c_int, c_ptr
      integer(c_int), value :: status
    end function psim_status_string_ptr
  end interface
end module psim_binding
