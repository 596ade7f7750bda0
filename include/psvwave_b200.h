/* psvwave_b200 -- C ABI of the B200-native P-SV velocity-stress solver.
 *
 * This is the drop-in boundary for psvWave's solver core: every entry point below replaces
 * one public member of the reference's C++ class `fdModel` (src/fdModel.h:35-318 of
 * larsgeb/psvWave) or one accessor of its pybind11 wrapper (src/psvWave.cpp:85-458).  The
 * reference-side binding a maintainer would add is shown in INTEGRATION.md; the pybind11
 * module shipped here (psvwave_b200/csrc/pybind_module.cpp) is that binding.
 *
 * Conventions
 *  - one opaque handle per fdModel object; a handle lives on ONE CUDA device (the device
 *    that is current, or set with psv_set_default_device, when it is created);
 *  - every call returns a status (0 = ok); the message of the last failure on the calling
 *    thread is returned by psv_last_error();
 *  - all pointers are HOST pointers owned by the caller unless the name says `device`;
 *    getters copy out, setters copy in (the reference's getters return copies too,
 *    src/psvWave.cpp:248-322);
 *  - calls are synchronous from the caller's point of view (exception: the simulate calls of an
 *    object that psv_set_devices spread over several GPUs, see there);
 *  - arrays are C-contiguous float64 with the reference's shapes: grid (nx, nz) with iz
 *    fastest (src/contiguous_arrays.cpp:14-21), seismograms (n_shots, nr, nt), snapshots
 *    (n_shots, snapshots, nx, nz);
 *  - there is no CPU fallback: without a usable CUDA device creation fails with
 *    PSV_ERR_CUDA.
 */
#ifndef PSVWAVE_B200_H
#define PSVWAVE_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct psv_model *psv_handle;

/* ---- status codes (mapped back to the reference's exception types by the binding) ---- */
#define PSV_OK 0
#define PSV_ERR_INVALID_ARGUMENT 1 /* std::invalid_argument  -> Python ValueError */
#define PSV_ERR_LENGTH 2           /* std::length_error      -> Python ValueError */
#define PSV_ERR_CUDA 3             /* CUDA runtime failure   -> Python RuntimeError */
#define PSV_ERR_STATE 4            /* call order violated (e.g. adjoint without stored forward) */
#define PSV_ERR_IO 5               /* file could not be written */
#define PSV_ERR_NCCL 6             /* NCCL could not be loaded / a collective failed -> Python RuntimeError */

const char *psv_last_error(void);
const char *psv_version(void);

/* ---- construction ------------------------------------------------------------------------
 * psv_config mirrors the reference's explicit 30-argument constructor
 * (src/fdModel.h:50-64; fdModel.cpp:29-62).  which_source_to_fire_in_which_shot is passed in
 * CSR form: sources of shot s are shot_sources[shot_offsets[s] .. shot_offsets[s+1]).
 * Source/receiver indices are relative to the physical domain (np_boundary is added by the
 * library, src/fdModel.cpp:197-212). */
typedef struct psv_config {
  int nt, nx_inner, nz_inner, nx_inner_boundary, nz_inner_boundary;
  double dx, dz, dt;
  int np_boundary;
  double np_factor;
  double scalar_rho, scalar_vp, scalar_vs;
  int npx, npz;
  double peak_frequency, source_timeshift, delay_cycles_per_shot;
  int n_sources, n_shots;
  const int *ix_sources, *iz_sources;  /* [n_sources] */
  const double *moment_angles;         /* [n_sources], degrees */
  const int *shot_offsets;             /* [n_shots + 1] */
  const int *shot_sources;             /* [shot_offsets[n_shots]] */
  int nr;
  const int *ix_receivers, *iz_receivers; /* [nr] */
  int snapshot_interval;
  const char *observed_data_folder, *stf_folder;
} psv_config;

/* fdModel::fdModel(const char*)  src/fdModel.cpp:16-27, .ini keys src/fdModel.cpp:470-523 */
int psv_create_from_ini(const char *ini_path, psv_handle *out);
/* fdModel::fdModel(30 args)      src/fdModel.cpp:29-62 */
int psv_create(const psv_config *config, psv_handle *out);
/* fdModel::fdModel(const fdModel&) src/fdModel.cpp:64-97,301-388 (deep copy incl. snapshots) */
int psv_clone(psv_handle src, psv_handle *out);
/* fdModel::~fdModel              src/fdModel.cpp:99-137 */
int psv_destroy(psv_handle h);

/* Host-only dry run of psv_create_from_ini: read + validate the .ini exactly as construction
 * would (same status codes / messages) without touching a device, and write the derived
 * configuration as "key=value" lines (lists as a,b,c; shots separated by '|') into `summary`
 * (truncated to `capacity`).  Needs no GPU. */
int psv_check_ini(const char *ini_path, char *summary, size_t capacity);

/* CUDA device used by subsequent psv_create* calls on this thread (-1 = current device). */
int psv_set_default_device(int device);
int psv_device_count(int *count);

/* ---- several GPUs behind ONE object (no counterpart in the reference, whose
 * parallelism is OpenMP inside each call) ------------------------------------------------------
 * psv_set_devices(h, ids, n): the object spans the n CUDA devices ids[0..n) of this box.  The
 * medium, the observed data and all options are replicated; shot s belongs to ids[s mod n] (a
 * fixed map).  From then on psv_forward_simulate / psv_adjoint_simulate only QUEUE the run on the
 * shot's GPU (one worker thread per GPU inside the handle) and return, so the reference's own
 * user loop   for shot: forward_simulate(shot) ... for shot: adjoint_simulate(shot)
 * (docs-source/source/typical-workflow.rst:5-13,42-59; fdModel::run_model,
 * src/fdModel.cpp:1291-1308) keeps all GPUs busy; every call that reads a result waits for what it
 * needs.  A failure inside a queued run is reported by the next call that waits.  Synthetics are
 * gathered onto ids[0] device to device before the misfit (summed there, in the reference's
 * order); the per-GPU lambda/mu/rho accumulations (src/fdModel.cpp:832-858) are summed by ONE
 * ncclAllReduce(sum, double) over the contiguous kernel buffer, issued by whatever reads the
 * kernels first (psv_map_kernels_to_velocity, psv_get_array of a Lame kernel, psv_write_kernels):
 * native NCCL inside the library, no PyTorch, no launcher.  If ids[0] is not the device the object
 * lives on, the object moves there.  n = 1 returns to a single device (stored snapshots of shots
 * that ran elsewhere are dropped).  Monte-Carlo batches and the domain split need single-device
 * handles.  The environment variable PSVWAVE_DEVICES ("all" or "0,1,...") applies the same call
 * to every object at creation, for scripts that cannot be edited at all. */
int psv_set_devices(psv_handle h, const int *device_ids, int n);
/* *n = number of devices of the object; the first min(*n, capacity) ids are written to device_ids. */
int psv_get_devices(psv_handle h, int *device_ids, int capacity, int *n);

/* ---- the hot path ---------------------------------------------------------------------- */
/* fdModel::forward_simulate       src/fdModel.cpp:526-787 (binding: src/psvWave.cpp:112-158) */
int psv_forward_simulate(psv_handle h, int i_shot, int store_fields, int verbose,
                         int output_wavefields);
/* fdModel::adjoint_simulate       src/fdModel.cpp:789-969 (binding: src/psvWave.cpp:160-205) */
int psv_adjoint_simulate(psv_handle h, int i_shot, int verbose);
/* fdModel::calculate_l2_misfit    src/fdModel.cpp:1128-1145; also stored, see PSV_REAL_MISFIT */
int psv_calculate_l2_misfit(psv_handle h, double *misfit_out);
/* fdModel::calculate_l2_adjoint_sources src/fdModel.cpp:1147-1162 */
int psv_calculate_l2_adjoint_sources(psv_handle h);
/* fdModel::reset_kernels          src/fdModel.cpp:1310-1324 */
int psv_reset_kernels(psv_handle h);
/* fdModel::map_kernels_to_velocity src/fdModel.cpp:1164-1183 */
int psv_map_kernels_to_velocity(psv_handle h);
/* fdModel::update_from_velocity   src/fdModel.cpp:1034-1051 */
int psv_update_from_velocity(psv_handle h);
/* fdModel::run_model              src/fdModel.cpp:1291-1308 */
int psv_run_model(psv_handle h, int verbose, int simulate_adjoint);

/* ---- model / gradient vectors (src/fdModel.cpp:1358-1515) ------------------------------- */
int psv_get_model_vector(psv_handle h, double *out, size_t n);
int psv_set_model_vector(psv_handle h, const double *in, size_t n);
int psv_get_gradient_vector(psv_handle h, double *out, size_t n);

/* ---- scalar attributes (read-only attributes of the Python class, src/psvWave.cpp:499-626) */
typedef enum psv_int_key {
  PSV_INT_NT = 0, PSV_INT_NX, PSV_INT_NZ, PSV_INT_NX_INNER, PSV_INT_NZ_INNER,
  PSV_INT_NX_INNER_BOUNDARY, PSV_INT_NZ_INNER_BOUNDARY, PSV_INT_NP_BOUNDARY,
  PSV_INT_N_SOURCES, PSV_INT_N_SHOTS, PSV_INT_NR, PSV_INT_SNAPSHOT_INTERVAL,
  PSV_INT_SNAPSHOTS, PSV_INT_NX_FREE_PARAMETERS, PSV_INT_NZ_FREE_PARAMETERS,
  PSV_INT_FREE_PARAMETERS, PSV_INT_BASIS_GRIDPOINTS_X, PSV_INT_BASIS_GRIDPOINTS_Z,
  PSV_INT_DEVICE, PSV_INT_N_DEVICES, PSV_INT_KEY_COUNT_
} psv_int_key;
typedef enum psv_real_key {
  PSV_REAL_DX = 0, PSV_REAL_DZ, PSV_REAL_DT, PSV_REAL_NP_FACTOR, PSV_REAL_SCALAR_RHO,
  PSV_REAL_SCALAR_VP, PSV_REAL_SCALAR_VS, PSV_REAL_PEAK_FREQUENCY, PSV_REAL_ALPHA,
  PSV_REAL_T0, PSV_REAL_DELAY_CYCLES_PER_SHOT, PSV_REAL_MISFIT, PSV_REAL_KEY_COUNT_
} psv_real_key;
int psv_get_int(psv_handle h, psv_int_key key, long long *out);
int psv_get_real(psv_handle h, psv_real_key key, double *out);

/* ---- arrays ------------------------------------------------------------------------------
 * Names follow the reference's public fields (src/fdModel.h:209-252). */
typedef enum psv_array {
  /* grid arrays, (nx, nz) */
  PSV_VX = 0, PSV_VZ, PSV_TXX, PSV_TZZ, PSV_TXZ,
  PSV_LM, PSV_LA, PSV_MU, PSV_B_VX, PSV_B_VZ, PSV_RHO, PSV_VP, PSV_VS,
  PSV_LAMBDA_KERNEL, PSV_MU_KERNEL, PSV_DENSITY_L_KERNEL,
  PSV_VP_KERNEL, PSV_VS_KERNEL, PSV_DENSITY_V_KERNEL, PSV_TAPER,
  /* signals */
  PSV_T,      /* (nt) */
  PSV_STF,    /* (n_sources, nt) */
  PSV_MOMENT, /* (n_sources, 2, 2) */
  /* seismograms, (n_shots, nr, nt) */
  PSV_RTF_UX, PSV_RTF_UZ, PSV_RTF_UX_TRUE, PSV_RTF_UZ_TRUE, PSV_A_STF_UX, PSV_A_STF_UZ,
  /* forward snapshots, (n_shots, snapshots, nx, nz); shots never stored read as zeros */
  PSV_ACCU_VX, PSV_ACCU_VZ, PSV_ACCU_TXX, PSV_ACCU_TZZ, PSV_ACCU_TXZ,
  PSV_ARRAY_COUNT_
} psv_array;

int psv_array_size(psv_handle h, psv_array which, size_t *count);
/* Copy out `count` doubles (must equal psv_array_size). */
int psv_get_array(psv_handle h, psv_array which, double *out, size_t count);
/* Copy in.  Settable: PSV_VP/PSV_VS/PSV_RHO (no implicit update_from_velocity),
 * PSV_RTF_UX/UZ, PSV_RTF_UX_TRUE/UZ_TRUE, PSV_A_STF_UX/UZ, PSV_STF, PSV_MOMENT,
 * PSV_LAMBDA_KERNEL/PSV_MU_KERNEL/PSV_DENSITY_L_KERNEL. */
int psv_set_array(psv_handle h, psv_array which, const double *in, size_t count);
/* fdModelExtended::set_parameter_fields (src/psvWave.cpp:256-297): copy + update_from_velocity */
int psv_set_parameter_fields(psv_handle h, const double *vp, const double *vs, const double *rho,
                             size_t count);

typedef enum psv_index_list {
  PSV_IX_SOURCES = 0, PSV_IZ_SOURCES, PSV_IX_RECEIVERS, PSV_IZ_RECEIVERS, /* incl. np_boundary */
  PSV_SHOT_OFFSETS, PSV_SHOT_SOURCES
} psv_index_list;
int psv_index_list_size(psv_handle h, psv_index_list which, size_t *count);
int psv_get_index_list(psv_handle h, psv_index_list which, int *out, size_t count);

/* ---- text I/O next to the path (same file formats as the reference) ---------------------- */
/* fdModel::write_receivers(prefix) src/fdModel.cpp:971-1006 */
int psv_write_receivers(psv_handle h);
/* fdModel::load_receivers          src/fdModel.cpp:1053-1126 */
int psv_load_receivers(psv_handle h, int verbose);
/* fdModel::write_sources           src/fdModel.cpp:1008-1032 */
int psv_write_sources(psv_handle h);
/* fdModel::write_kernels           src/fdModel.cpp:1326-1356 */
int psv_write_kernels(psv_handle h);
/* fdModel::load_model              src/fdModel.cpp:1185-1289 */
int psv_load_model(psv_handle h, const char *de_path, const char *vp_path, const char *vs_path,
                   int verbose);
/* fdModel::load_vector             src/fdModel.cpp:1517-1572; n must be 3*nx_free*nz_free */
int psv_load_vector(psv_handle h, const char *path, int verbose, double *out, size_t n);

/* ---- additive entry points (not in the reference) ----------------------------------------- */
/* Drop the stored forward snapshots of one shot (i_shot >= 0): its buffer goes back to the
 * model's pool and is reused by the next stored forward run.  i_shot = -1 drops all shots and
 * frees the pool.  Lets a per-shot forward->adjoint loop run Marmousi-scale problems whose
 * all-shots snapshot set (src/fdModel.cpp:185-190) would not fit in HBM. */
int psv_release_snapshots(psv_handle h, int i_shot);
/* L2 misfit of ONE shot (the terms of src/fdModel.cpp:1140-1141 for that shot only); does not
 * touch the stored misfit.  Ranks that own a subset of the shots sum these. */
int psv_calculate_l2_misfit_shot(psv_handle h, int i_shot, double *misfit_out);
/* Device pointer + element count of the three Lame-basis kernels stored back to back
 * [lambda | mu | density_l], each nx*ld doubles with row pitch ld (see PSV_TUNE "pitch");
 * used by the host layer for the cross-rank sum (one all-reduce) over NCCL. */
int psv_device_kernel_buffer(psv_handle h, void **device_ptr, size_t *count);
int psv_synchronize(psv_handle h);
/* Gradient of all shots, one shot at a time: forward(store) -> misfit of the shot ->
 * adjoint sources -> adjoint (-> psv_release_snapshots of that shot if release_each_shot), then
 * psv_map_kernels_to_velocity.  Same result as psv_run_model(h, 0, 1) (src/fdModel.cpp:1291-1308)
 * with one shot's snapshots resident instead of all; *misfit_out = sum of the per-shot misfits. */
int psv_run_gradient(psv_handle h, int release_each_shot, double *misfit_out);
/* Device-resident steepest descent (the loop of the reference's
 * tests/python/test_descent.py:106-129): model vector <- model vector - step_length * gradient
 * vector on the free-window basis, then update_from_velocity; bit-identical to
 * get_model_vector / get_gradient_vector / set_model_vector on the host.  psv_gradient_descent reads
 * the model vector ONCE into device memory and then runs `iterations` rounds of { m -= step * g;
 * set_model_vector(m); psv_run_gradient } like the reference loop (which never re-reads m), reporting
 * the misfit after each round (misfits_out may be NULL); the kernels of the starting model must
 * exist (psv_run_gradient). */
int psv_descent_step(psv_handle h, double step_length);
int psv_gradient_descent(psv_handle h, int iterations, double step_length, int release_each_shot,
                         double *misfits_out);
/* The rest of the reference's outer loop on the device.  The reference's notebooks and docs
 * (docs-source/source/typical-workflow.rst:16-72; Notebook 2 cells 15-17) hand get_model_vector / get_gradient_vector to
 * SciPy: scipy.ndimage.gaussian_filter(panel.reshape(nz_free, nx_free), sigma) on the three thirds of the gradient vector,
 * then scipy.optimize.minimize(method="L-BFGS-B", maxcor=10, maxiter=50, gtol=1e-6).
 * psv_get_gradient_vector_smoothed: the gradient vector after exactly that filter (separable, mode "reflect", truncate 4.0,
 * scipy's summation order), computed on the device; sigma = 0 is psv_get_gradient_vector.
 * psv_lbfgs: L-BFGS with `history` (s, y) pairs on the device-resident model vector -- two-loop recursion, initial scaling
 * s.y / y.y, first step 1/|g| like L-BFGS-B, unit step afterwards with backtracking on the sufficient-decrease
 * condition -- every evaluation being psv_run_gradient (+ the smoothing when smoothing_sigma > 0).  Starts from the
 * object's current medium, stops after max_iterations or when max|g| <= gtol, leaves the final medium in the object.
 * misfits_out[0 .. *iterations_done] (capacity max_iterations + 1, may be NULL) receives the misfit before the first and
 * after every iteration.  End state: medium, synthetics, sensitivity kernels and misfit all belong to the returned point
 * (when the line search gives up, that point is evaluated once more; it counts in *evaluations_done).  Smoothing calls on
 * one GPU are serialised inside the library (one set of filter weights per device).  Not bit-comparable with SciPy's L-BFGS-B (another line search); the tests compare the misfit
 * reached on the notebook-2 problem. */
int psv_get_gradient_vector_smoothed(psv_handle h, double sigma, double *out, size_t n);
/* Host only (no device needed): the 2 * radius + 1 filter weights psv_get_gradient_vector_smoothed uses for `sigma` --
 * scipy.ndimage's _gaussian_kernel1d(sigma, 0, int(4 sigma + 0.5)) -- written to weights_out (capacity doubles). */
int psv_gaussian_weights(double sigma, double *weights_out, size_t capacity, int *radius_out);
int psv_lbfgs(psv_handle h, int max_iterations, int history, double gtol, double smoothing_sigma,
              int release_each_shot, double *misfits_out, int *iterations_done, int *evaluations_done);
/* Monte-Carlo ensembles (SURVEY 8e "replicas only"; the reference's route is n independent copies of
 * the object, src/fdModel.cpp:64-97, each simulated on its own).  The n models must describe one
 * experiment on one device (same grid, time axis, boundary, sources, receivers, snapshot interval);
 * media differ.  All n models advance through forward_simulate / adjoint_simulate of shot i_shot
 * together, one kernel launch per time step for the whole ensemble, and end in exactly the state n
 * separate psv_forward_simulate / psv_adjoint_simulate calls would leave.  Tuning options are
 * those of handles[0].  Needs a fused variant (>= 2, the default). */
int psv_forward_batch(const psv_handle *handles, int n, int i_shot, int store_fields);
int psv_adjoint_batch(const psv_handle *handles, int n, int i_shot);
/* ---- domain split for grids too large for one GPU (no counterpart in the reference) --
 * The grid is cut into `nranks` x-slabs (rows are contiguous in z, so a halo is a block of whole
 * rows).  A slab handle is an fdModel over the rows it owns plus `halo_rows` rows of each
 * neighbour: grid arrays (get/set) have the LOCAL shape (nx_local, nz), sources/receivers/time
 * axis/seismograms are those of the whole experiment.  One time step consumes 4 halo rows, so
 * halo_rows = 4k lets k steps run between two exchanges.  The host layer drives all slabs in
 * lock-step:   begin;  repeat { steps(k); pack; <swap send/recv buffers with the neighbours>;
 * unpack; }  end;   seismograms are valid for receivers on owned rows, kernels on owned rows.
 * psv_forward_simulate / psv_adjoint_simulate refuse a slab that has neighbours. */
int psv_create_slab_from_ini(const char *ini_path, int rank, int nranks, int halo_rows, psv_handle *out);
/* out[7] = nx_global, first global row (x offset), nx_local, halo rows below, halo rows above,
 * row pitch in doubles, 1 if the handle is a slab */
int psv_slab_info(psv_handle h, long long out[7]);
int psv_march_begin(psv_handle h, int i_shot, int adjoint, int store_fields);
/* Enqueue up to n time steps on the model's stream (no wait); *done = steps enqueued, 0 at the end. */
int psv_march_steps(psv_handle h, int n, int *done);
int psv_march_end(psv_handle h);
/* All slabs of one grid inside ONE process -- one per GPU of the box (create each with psv_set_default_device(d)
 * first), or several on one GPU -- march through forward_simulate (adjoint = 0) or adjoint_simulate of shot i_shot
 * together.  Every halo_rows/4 steps each slab pulls the edge rows its neighbours own straight out of their state
 * arrays into its own halo rows: peer-to-peer loads over NVLink by one copy kernel per slab, ordered between the
 * slabs' streams by CUDA events; no staging buffers, no NCCL, no host synchronisation inside the march.  `slabs`
 * holds ALL slabs in slab order.  Results are those of psv_march_begin / steps / pack-exchange-unpack / end. */
int psv_march_slabs(const psv_handle *slabs, int n, int i_shot, int adjoint, int store_fields);
/* Stage the owned edge rows of vx vz txx tzz txz into the send buffers / the received rows into
 * the halo (stream-ordered kernels).  Buffers: [5 fields][halo_rows * pitch] doubles per side,
 * `count` doubles each; a side without neighbour has NULL pointers. */
int psv_halo_pack(psv_handle h);
int psv_halo_unpack(psv_handle h);
int psv_halo_buffers(psv_handle h, void **send_lo, void **send_hi, void **recv_lo, void **recv_hi,
                     size_t *count);
/* The CUDA stream (cudaStream_t) every kernel and copy of this model is ordered on. */
int psv_stream(psv_handle h, void **stream);
/* Device stopwatch on the model's stream (CUDA events): start records an event, stop records a
 * second one, waits for it and returns the milliseconds between the two. */
int psv_timer_start(psv_handle h);
int psv_timer_stop(psv_handle h, double *milliseconds);

/* Page-locked host memory for callers that want the library's copies to run at full PCIe / C2C speed (the shipped
 * binding backs the large arrays its getters return with such buffers and recycles them): plain cudaHostAlloc /
 * cudaFreeHost, no handle needed.  psv_host_alloc fails with PSV_ERR_CUDA / PSV_ERR_LENGTH when the memory cannot be
 * pinned; any host pointer is accepted by every getter and setter. */
int psv_host_alloc(size_t bytes, void **host_ptr);
int psv_host_free(void *host_ptr);

/* The L2 misfit of psv_calculate_l2_misfit, computed without allocating anything.  NO COUNTERPART IN THE REFERENCE:
 * it exists for callers that may not reach cudaMalloc -- a host program that gave all GPU memory to its own arena
 * and forbids plug-ins to ask the driver for more, or a sampling loop that cannot afford the device-wide stall every
 * cudaMalloc / cudaFree brings.  Everything is DEVICE memory owned by the caller: the four seismogram arrays
 * (synthetic ux, uz; observed ux, uz; n doubles each, any contiguous range in (shot, receiver, time) order) and the
 * one double the result is written to.  The block partial sums, which the handle-based call keeps in a cudaMalloc'ed
 * buffer, live in a fixed 64 KiB array inside the library image instead; each GPU that loads the library holds that
 * array until the process ends.
 *  - limit: n <= PSV_MISFIT_STATIC_MAX_SAMPLES per array, larger n fails with PSV_ERR_LENGTH and launches nothing
 *    (sum several ranges yourself, or use the handle-based call);
 *  - runs on the device that is current in the calling thread, ordered on `stream` (a cudaStream_t, NULL = default
 *    stream); returns once the two kernels are enqueued, the result is valid after the stream reaches that point;
 *  - the first call on each device looks up the array's device address (cudaGetSymbolAddress); make that call before
 *    any stream capture begins;
 *  - the array is shared by every call on that GPU and keeps the previous call's values (harmless: each call writes
 *    the slots it later reads); calls that overlap in time on one GPU race on it -- keep them on one stream;
 *  - value: bit-identical to psv_calculate_l2_misfit / psv_calculate_l2_misfit_shot on the same arrays with
 *    dt = PSV_REAL_DT (same terms 0.5*dt*d^2, same order of additions). */
#define PSV_MISFIT_STATIC_MAX_SAMPLES 33554432 /* 8192 block partials x 4096 samples per block */
int psv_l2_misfit_static_scratch(const double *device_ux, const double *device_uz, const double *device_obs_ux,
                                 const double *device_obs_uz, size_t n, double dt, double *device_misfit_out,
                                 void *stream);

/* Tuning / instrumentation knobs, by name.  Unknown names fail with PSV_ERR_INVALID_ARGUMENT.
 *   "variant"  which step kernels run the time loop (results are the same):
 *              0 one thread per point, two sweeps      1 x-marching two sweeps
 *              2 one fused launch per time step        3 (default) the same, fed by a TMA bulk-copy ring
 *              4 two time steps per launch (temporal blocking experiment)
 *   "resident" the WHOLE time loop in one cooperative launch for grids that fit in the SMs' shared memory
 *              (resident.cuh): -1 (default) = whenever the grid fits, 0 = never, 1 = required;
 *              "resident_rows" rows per CTA (0 = automatic)
 *   "exact"    1 = IEEE operation order of the reference (true divisions, no FMA): results are
 *              bit-identical to the reference build; 0 (default) = reciprocals + FMA (<= 1e-10)
 *   "fp32"     1 = optional single-precision wavefield mode (state, medium and snapshots in float,
 *              seismograms and sensitivity kernels in double); stated tolerance 1e-3 relative L2
 *   "profile"  1 = record CUDA events around sampled step launches (see psv_get_stat)
 *   "graph"    the time loop of one (direction, shot, store) captured into a CUDA graph on first use and replayed
 *              afterwards: 1 = always, 0 = never, -1 (default) = when a step launch covers at most 2^20 grid points
 *              (the step kernel is then shorter than the gap between two stream launches) and the same loop is
 *              run a second time; "graphs_cached" reads
 *              the number of graphs held.  "pdl" 1 = programmatic dependent launch of consecutive variant-3 steps
 *   "tune"     launch geometry of the default step kernel chosen by MEASUREMENT at the first time loop of a model (a few
 *              milliseconds of plain step launches over its own state arrays): -1 (default) = grids of >= 2^21 points,
 *              0 = never (formula), 1 = always; "tuned_warps" / "tuned_strip" / "tune_ms" read the outcome (psv_get_stat)
 *   launch geometry (sweeps under profiles/): "strip" "block" (variants 0/1), "fused_strip"
 *              "fused_block" (2), "pipe_warps" "pipe_stages" "l2_mode" "l2_persist_mb" (3),
 *              "pair_warps" "pair_stages" "pair_strip" (4) */
int psv_set_option(psv_handle h, const char *name, long long value);
/* Number of sweep-kernel variants compiled into the library ("variant" < this). */
int psv_variant_count(void);
int psv_get_option(psv_handle h, const char *name, long long *value);
/* Counters of the last forward/adjoint call:
 *   "elapsed_ms" device time of the whole time loop (CUDA events on the model's stream)
 *   "launches"   kernels launched by that call
 *   "stress_ms", "velocity_ms"  mean duration of one sampled launch (needs "profile" = 1)
 *   "pitch"      row pitch (in doubles) of the device grid arrays
 *   "resident_ctas" CTAs of the resident whole-loop kernel in that call (0: per-step launches)
 *   "devices", "allreduces", "allreduce_ms"  of an object on several GPUs (psv_set_devices) */
int psv_get_stat(psv_handle h, const char *name, double *value);

#ifdef __cplusplus
}
#endif
#endif /* PSVWAVE_B200_H */
