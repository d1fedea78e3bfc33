/* gapflow_b200 -- C ABI of the B200-native GaPFlow hot path.
 *
 * GaPFlow (hannes-holey/GaPFlow) has no FFI or plugin interface: its hot path is Python
 * methods on NumPy arrays.  Each entry point below states the reference interface it
 * replaces (file:line relative to the GaPFlow repository).  INTEGRATION.md shows the
 * ctypes binding a maintainer would add to the reference.
 *
 * Conventions
 *  - plain C, no C++/torch types; every function returns 0 (GPF_OK) or a negative
 *    GPF_E_* code; gpf_last_error() gives the text for the calling thread.
 *  - "dense" arrays have the reference's layout: float64, C order,
 *    shape (ncomp, Nx+2, Ny+2) (one ghost layer, GaPFlow/problem.py:122-124).
 *  - `stream` arguments are cudaStream_t passed as void* (0 = legacy default stream).
 *    Functions enqueue work on that stream and do not synchronise unless they return
 *    host scalars (documented per function).
 *  - a handle is bound to the CUDA device current at creation; not thread-safe.
 */
#ifndef GAPFLOW_B200_H
#define GAPFLOW_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPF_OK 0
#define GPF_E_INVALID (-1)     /* bad argument / unsupported configuration */
#define GPF_E_CUDA (-2)        /* CUDA runtime or driver error             */
#define GPF_E_NOMEM (-3)
#define GPF_E_STATE (-4)       /* call sequence error                       */

/* equation of state ids (GaPFlow/models/pressure.py:49-70) */
enum { GPF_EOS_DH = 0, GPF_EOS_PL = 1, GPF_EOS_VDW = 2, GPF_EOS_MT = 3,
       GPF_EOS_CUBIC = 4, GPF_EOS_BWR = 5, GPF_EOS_BAYADA = 6 };

/* boundary kinds per side and component */
enum { GPF_BC_P = 0, GPF_BC_D = 1, GPF_BC_N = 2, GPF_BC_HALO = 3 };
/* sides: x0 = array index 0, x1 = index Nx+1, y0 = index 0, y1 = index Ny+1 */
enum { GPF_SIDE_X0 = 0, GPF_SIDE_X1 = 1, GPF_SIDE_Y0 = 2, GPF_SIDE_Y1 = 3 };

/* arithmetic contract of the pass kernels (gpf_fv_cfg.arith) */
enum { GPF_ARITH_EXACT = 0, GPF_ARITH_FAST = 1 };

/* viscosity modifiers (GaPFlow/models/viscosity.py:34-96): piezoviscosity and shear-thinning law ids.
 * GPF_*_IDENTITY is a block whose name the reference does not know: the law is the identity
 * (viscosity.py:64, 94) but the viscosity is still evaluated as a field. */
enum { GPF_PIEZO_ABSENT = 0, GPF_PIEZO_IDENTITY = 1, GPF_PIEZO_BARUS = 2, GPF_PIEZO_ROELANDS = 3,
       GPF_PIEZO_DUKLER = 4, GPF_PIEZO_MCADAMS = 5 };
enum { GPF_THIN_ABSENT = 0, GPF_THIN_IDENTITY = 1, GPF_THIN_EYRING = 2, GPF_THIN_CARREAU = 3 };

/* field ids for upload/download */
enum { GPF_F_Q = 0, GPF_F_TOPO = 1, GPF_F_EXTRA = 2 };

/* Problem description: the sanitised `grid`, `geometry`, `numerics`, `properties`
 * dicts of GaPFlow/io.py:139-395, flattened. */
typedef struct gpf_fv_cfg {
    int32_t Nx, Ny;            /* interior cells of THIS rank's slab */
    double dx, dy;
    /* Boundary table.  Build it with gpf_fv_cfg_set_bc() from the reference's 12 mask
     * lists; it encodes the reference's cross-lookups (GaPFlow/problem.py:676-768). */
    int32_t bc_type[4][3];     /* GPF_BC_* per side, component            */
    int32_t bc_src[4][3];      /* component whose adjacent value is used  */
    double bc_val[4];          /* Dirichlet target per side               */
    double U, V;               /* wall velocities (geometry)              */
    double eta, zeta;          /* shear, bulk viscosity (properties)      */
    int32_t eos;               /* GPF_EOS_*                               */
    double eos_par[8];         /* DH: rho0,P0,C1,C2 | PL: rho0,P0,alpha | vdW: M,T,a,b |
                                  MT: rho0,P0,K,n | cubic: a,b,c,d | BWR: T,gamma |
                                  Bayada: rho_l,rho_v,c_l,c_v              */
    double rho0;               /* initial density (properties['rho0'])    */
    int32_t adaptive;          /* numerics['adaptive']                    */
    double CFL, dt, tol;       /* numerics                                */
    int32_t max_it, mc_order;  /* numerics['max_it'], numerics['MC_order'] */
    int32_t has_extra;         /* 0: slip-length field identically zero   */
    int32_t halo_edge[2];      /* GPF_BC_HALO x side that is the global domain's periodic edge: its
                                  ghost row is part of the reference's padded array and is counted
                                  in the reductions (problem.py:337-347) by this rank */
    int32_t arith;             /* GPF_ARITH_EXACT (0): every operation rounded as the reference's NumPy
                                  expressions round it -> bit-identical fields for the + - * / sqrt EOS;
                                  GPF_ARITH_FAST (1): FMA contraction and a*rcp(b) quotients in the pass
                                  kernels -> within 1e-12 relative per step, ~1.4x faster */
    int32_t visc_field;        /* != 0: properties has a 'piezo' and / or 'thinning' block, the shear viscosity is
                                  a per-cell field (stress.py:305-326); gpf_fv_set_viscosity() must follow
                                  gpf_fv_create().  Adds one plane + the gradient tables to the workspace */
    int32_t reserved[3];
} gpf_fv_cfg;

/* Host-visible scalars of the time loop (GaPFlow/problem.py:314-362, 571-586). */
typedef struct gpf_fv_scalars {
    int64_t step;              /* Problem.step                            */
    double simtime, dt;        /* Problem.simtime, Problem.dt             */
    double residual;           /* Problem.residual                        */
    double residual_buffer[5]; /* Problem.residual_buffer (oldest first)  */
    int32_t residual_count;    /* entries valid in residual_buffer        */
    double ekin, ekin_old;     /* kinetic_energy (of current q), kinetic_energy_old */
    double v_max, v_sound;     /* Problem.v_max, Problem.pressure.v_sound */
    double dt_crit, cfl;
    int32_t has_nan, has_neg;  /* q_has_nan, q_has_negative_density of the LAST attempted step */
    int32_t stop;              /* Problem._stop (set by _finalize)        */
    int32_t converged;         /* Problem.converged                       */
} gpf_fv_scalars;

typedef struct gpf_fv gpf_fv;  /* opaque solver handle */

const char* gpf_last_error(void);
int gpf_version(void);

/* Fill cfg->bc_* from the reference's boolean masks (io.py:176-217).
 * masks[s][k][c]: s in {xE,xW,yS,yN}, k in {P,D,N}, c component; d_val[s] = bc_<s>_D_val.
 * Reproduces GaPFlow/problem.py:676-768 including its E/W and S/N cross-lookups.
 * Returns GPF_E_INVALID for mask combinations on which the reference itself raises
 * (shape mismatch) or leaves ghost cells undefined (mixed P / non-P on one side). */
int gpf_fv_cfg_set_bc(gpf_fv_cfg* cfg, const uint8_t masks[4][3][3], const double d_val[4]);

/* Device memory needed by a solver for cfg (bytes); the caller allocates it (e.g. a
 * torch uint8 tensor) and keeps it alive for the handle's lifetime. */
int gpf_fv_workspace_bytes(const gpf_fv_cfg* cfg, size_t* bytes);

/* Replaces Problem.__init__/_initialize/_pre_run for the FV fields
 * (GaPFlow/problem.py:77-152, 412-443, 662-670).  `workspace` is device memory.
 * Fields start as _initialize leaves them; upload topography (and extra) next, then
 * call gpf_fv_pre_run(). */
int gpf_fv_create(const gpf_fv_cfg* cfg, void* workspace, size_t bytes, gpf_fv** out);
int gpf_fv_destroy(gpf_fv* s);

/* Viscosity modifiers of a solver created with cfg.visc_field != 0.  Replaces the piezoviscosity /
 * shear_rate_avg / shear_thinning_factor chain at the head of WallStress.update and BulkStress.update
 * (GaPFlow/models/stress.py:305-326, 429-450; laws in GaPFlow/models/viscosity.py:99-318): before every
 * pass the viscosity of each cell of the padded array is evaluated from the pass's own pressure field,
 *   mu0 = piezo(p)  (piezo(rho) for the Bayada-Chupin EOS, stress.py:307),
 *   eta = mu0 * thinning(shear_rate_avg(dp/dx, dp/dy, h, U, V, mu0), mu0),
 * with dp/dx, dp/dy = np.gradient(p, x, axis) over the padded array (stress.py:171-192: central in
 * the interior, one-sided at the two edges, uniform or non-uniform form chosen as NumPy chooses it
 * from the coordinates).  piezo_par: Barus {aB} | Roelands {mu_inf, p_ref, z} | Dukler, McAdams
 * {eta_v, rho_l, rho_v};  thin_par: Eyring {tauE} | Carreau {mu_inf, lam, a, N}.  x (Nx+2), y (Ny+2):
 * HOST arrays, the cell-centre coordinates topo.x[:, 0], topo.y[0, :] (topography.py:38-54); may be
 * NULL when thinning is absent.  Slab (GPF_BC_HALO) solvers are refused: the gradient at a halo row
 * needs a second halo row.  exp / log / pow / asinh / hypot make this path tolerance-based (1e-12 per
 * step) in both arithmetic contracts. */
int gpf_fv_set_viscosity(gpf_fv* s, int piezo, const double* piezo_par, int thinning, const double* thin_par,
                         const double* x, const double* y);
/* the viscosity field of the current q (which = 0) or of the last corrector pass's input (which = 1),
 * dense (Nx+2, Ny+2), device or host; synchronises for host output */
int gpf_fv_viscosity(gpf_fv* s, int which, double* eta, int on_device, void* stream);

/* Dense (ncomp, Nx+2, Ny+2) float64 <-> solver storage.  ncomp: Q 3, TOPO 3 (h, dh/dx,
 * dh/dy; a 4-component array may be passed with ncomp=4, comp 3 is ignored), EXTRA 1.
 * `on_device` != 0: `data` is a device pointer.  Host pointers are copied with
 * cudaMemcpyAsync on `stream` (pinned memory makes them asynchronous).
 * Replaces direct writes to Problem.q / topo.full / extra (problem.py:314-317). */
int gpf_fv_upload(gpf_fv* s, int field, const double* data, int ncomp, int on_device, void* stream);
int gpf_fv_download(gpf_fv* s, int field, double* data, int ncomp, int on_device, void* stream);

/* Problem._pre_run (problem.py:428-443): step=0, simtime=0, residual=1, residual
 * buffer, dt = CFL*dt_crit or numerics['dt'];  also (re)computes kinetic_energy_old as
 * _initialize does (problem.py:670) when reset_ekin_old != 0.  Synchronises. */
int gpf_fv_pre_run(gpf_fv* s, int reset_ekin_old, void* stream);

/* n x Problem.update() (problem.py:509-569) enqueued on `stream`, no synchronisation.
 * until_converged != 0 reproduces the loop condition of Problem.run (problem.py:401):
 * steps become no-ops on the device once converged, step >= max_it or _stop. */
int gpf_fv_step(gpf_fv* s, int n, int until_converged, void* stream);

/* Ghost-cell refresh of the current q: Problem._communicate_ghost_buffers
 * (problem.py:676-707). */
int gpf_fv_fill_ghosts(gpf_fv* s, void* stream);

/* Reads the loop scalars (synchronises `stream`).  ekin, v_max, v_sound, dt_crit, cfl
 * are those of the current q as the reference's properties would return them. */
int gpf_fv_get_scalars(gpf_fv* s, gpf_fv_scalars* out, void* stream);
/* Overwrites step/simtime/dt/residual(+buffer)/ekin_old/stop from `in` (host is
 * authoritative, e.g. after a test assigned Problem.dt). */
int gpf_fv_set_scalars(gpf_fv* s, const gpf_fv_scalars* in, void* stream);

/* Problem.mass (problem.py:335-337); synchronises. */
int gpf_fv_mass(gpf_fv* s, double* mass, void* stream);

/* Closure fields of the current q as sol.nc stores them: pressure (1), wall_stress_xz
 * (12), wall_stress_yz (12), bulk_viscous_stress (3); dense outputs, device or host.
 * Replaces Pressure.update / WallStress.update / BulkStress.update
 * (GaPFlow/models/stress.py:289-362, 427-459, 600-622).  NULL outputs are skipped. */
int gpf_fv_closures(gpf_fv* s, double* pressure, double* wall_xz, double* wall_yz,
                    double* bulk, int on_device, void* stream);

/* Same, selecting the state the fields are evaluated on: which = 0 the current q; which = 1 the
 * input of the last corrector pass (the predictor's output), which is what the reference's
 * pressure / wall_stress / bulk fields hold when Problem.write() stores them: they are refreshed
 * at the START of each pass only (problem.py:532-540 vs :563, :629).  Valid after >= 1 step. */
int gpf_fv_closures_of(gpf_fv* s, int which, double* pressure, double* wall_xz, double* wall_yz,
                       double* bulk, int on_device, void* stream);

/* One-call host-buffer step for reference-side bindings: uploads q_host (dense, host),
 * runs n x update(), downloads q back, returns the scalars.  Synchronises.  For n == 1 (and
 * MC_order = +1, no viscosity field / surrogate / slab) upload, passes and download are pipelined
 * over chunks of tile rows on two internal copy streams; same result bit for bit.  Pinned host
 * memory makes the copies asynchronous (pageable memory works, staged by the driver). */
int gpf_fv_update_host(gpf_fv* s, double* q_host, int n, gpf_fv_scalars* out, void* stream);

/* --- slab decomposition along x (multi-GPU) ---------------------------------------
 * The reference runs on one process (problem.py:122-124; its _communicate_ghost_buffers,
 * problem.py:676-707, is a local copy), so these have no counterpart there.  A rank's
 * cfg has Nx = its slab's rows and GPF_BC_HALO on slab-internal x sides; gapflow_b200/dist.py
 * drives one Problem.update() as
 *   pass(0); halo_pack(1); <exchange>; halo_unpack(1);
 *   pass(1); halo_pack(2); <exchange>; halo_unpack(2);
 *   reduce_local(0); <all-gather of the 5-double records>; finish(GPF_FIN_STEP).
 * All calls enqueue on `stream` and never synchronise. */
enum { GPF_FIN_STEP = 0, GPF_FIN_PRERUN = 1, GPF_FIN_PRERUN_KEEP_EKIN = 2, GPF_FIN_REFRESH = 3 };
/* until_converged != 0: the steps driven through gpf_fv_pass / gpf_fv_finish / gpf_fv_finish_recv / gpf_fv_slab_step
 * honour the loop condition of Problem.run (problem.py:401) on the device, like gpf_fv_step(until_converged = 1):
 * once converged, at max_it or stopped they are no-ops (the exchanges still run, identically on every rank).
 * Set it identically on every rank. */
int gpf_fv_set_run_mode(gpf_fv* s, int until_converged);
/* one MacCormack pass (predictor: corrector = 0), problem.py:524-560 */
int gpf_fv_pass(gpf_fv* s, int corrector, void* stream);
/* the same pass in two launches, so that the exchange of the edge rows runs under the interior:
 *   pass_part(c, GPF_PART_EDGE, main); <event>; on a side stream: halo_pack, exchange, halo_unpack;
 *   pass_part(c, GPF_PART_INTERIOR, main); main waits for the side stream.
 * EDGE = the first and the last row of 8x64 tiles (rows 1..8 and the last <= 8 rows), INTERIOR =
 * the rest; EDGE + INTERIOR == ALL bit for bit (a pass reads only buffers it does not write). */
enum { GPF_PART_ALL = 0, GPF_PART_EDGE = 1, GPF_PART_INTERIOR = 2 };
int gpf_fv_pass_part(gpf_fv* s, int corrector, int part, void* stream);
/* rows 1 and Nx (y ghosts included) of buffer `rel` (0: current q, 1: predictor output, 2:
 * corrector output) -> sendbuf[2][3][Ny+2] (device);  recvbuf[0] -> row 0, recvbuf[1] -> row
 * Nx+1, HALO sides only. */
int gpf_fv_halo_pack(gpf_fv* s, int rel, double* sendbuf, void* stream);
int gpf_fv_halo_unpack(gpf_fv* s, int rel, const double* recvbuf, void* stream);
/* this rank's record {sum ekin, max v^2, max c^2, mass, flags} (5 doubles, device):
 * kind 0 folds the per-tile records of the last corrector pass, kind 1 reduces the current q
 * (halo rows excluded; they are counted by the rank that owns them) */
int gpf_fv_reduce_local(gpf_fv* s, int kind, double* rec5, void* stream);
/* advances the loop state from `nrec` gathered records (device, rank order):
 * _post_update / _finalize (GPF_FIN_STEP), _pre_run (GPF_FIN_PRERUN*), or only the scalars
 * of the current q (GPF_FIN_REFRESH) */
int gpf_fv_finish(gpf_fv* s, int mode, const double* recs, int nrec, void* stream);

/* --- GP surrogate closures in the FV step (GaPFlow/models/stress.py:346-356, 614-620) ---------
 * With a GP model the reference replaces, at the start of each pass, the pressure field by the GP
 * mean (Pressure.update) and the own shear component of both wall-stress fields by the two outputs
 * of the shear GP (WallStress.update: field[_out_index] = mean[0] lower wall, field[_out_index + 6] =
 * mean[1] upper wall; _out_index = 4 (xz model), 3 (yz model)); everything else stays fixed-form.
 * gpf_fv_set_surrogate hands the pass kernels the NORMALISED predictive means (what gpf_gp_mean /
 * gpf_gp_mean_var write) and the model's Yscale (gp.py:362: mean * Yscale): pressure (Np), shear_xz
 * and shear_yz (2, Np), device arrays over the padded grid in the row order of gpf_gp_features,
 * Np = (Nx+2)(Ny+2); NULL keeps the fixed-form law for that quantity; all NULL returns the solver
 * to its fixed-form kernels.  The arrays are read when a pass RUNS: the caller refreshes them from
 * the pass's input before each pass (rel = 0 predictor, 1 corrector in gpf_gp_features_of) and
 * drives one Problem.update() as
 *   features_of(0) -> gp_mean[_var] -> pass(0);  features_of(1) -> gp_mean -> pass(1);
 *   reduce_local(0, rec5);  features_of(2) -> gp_mean(..., &g);  finish_surrogate(GPF_FIN_STEP, rec5, c2)
 * gpf_fv_step is refused while surrogate arrays are set.  The surrogate passes run the general
 * kernel with plain IEEE divisions (tolerance contract of the GP path: 1e-9). */
int gpf_fv_set_surrogate(gpf_fv* s, const double* pressure, double pressure_scale,
                         const double* shear_xz, double shear_xz_scale,
                         const double* shear_yz, double shear_yz_scale);
/* gpf_fv_finish for `nrec` records (device, in/out; 1 for a single domain, the gathered records of all
 * ranks for slabs) after replacing their max c^2 by `c2` when a pressure GP is set: Pressure.v_sound's
 * GP branch (stress.py:522-537), c2 = max_i d mean / d x_0 * Yscale / X_scale[0] with the maximum from
 * gpf_gp_mean(..., dmean_dx0_max), over all ranks for slabs.  mode as gpf_fv_finish (GPF_FIN_PRERUN*
 * after reduce_local(1, rec5) reproduces _pre_run's dt = CFL * dt_crit).
 * Slabs: a rank's arrays cover ITS padded rows (halo rows included: they hold the neighbour's cells once
 * the exchange of the pass's input has run), so query cells are partitioned and the training set is
 * replicated; gapflow_b200/surrogate.py::SlabSurrogateDriver drives it. */
int gpf_fv_finish_surrogate(gpf_fv* s, int mode, double* recs, int nrec, double c2, void* stream);

/* --- pure functions (GaPFlow/models/*.py, GaPFlow/integrate.py); device pointers ---- */
/* eos_pressure(density, prop) -> pressure.py:35;  eos_sound_velocity -> sound.py:35 */
int gpf_eos_pressure(const gpf_fv_cfg* cfg, const double* dens, double* out, size_t n, void* stream);
int gpf_eos_sound_velocity(const gpf_fv_cfg* cfg, const double* dens, double* out, size_t n, void* stream);
/* stress_bottom / stress_top (6, n) and stress_avg (3, n), slip="top", dq=0
 * (viscous.py:37, 281, 612).  q (3, n), h (3, n), Ls (n) or NULL. */
int gpf_stress_bottom(const gpf_fv_cfg* cfg, const double* q, const double* h, const double* Ls,
                      double* out, size_t n, void* stream);
int gpf_stress_top(const gpf_fv_cfg* cfg, const double* q, const double* h, const double* Ls,
                   double* out, size_t n, void* stream);
int gpf_stress_avg(const gpf_fv_cfg* cfg, const double* q, const double* h, const double* Ls,
                   double* out, size_t n, void* stream);

/* The general signatures stress_bottom / stress_top / stress_avg(q, h, U, V, eta, zeta, Ls, dqx, dqy, slip)
 * (viscous.py:37, 281, 612): which = 0 bottom (6, n), 1 top (6, n), 2 avg (3, n); slip = GPF_SLIP_TOP (the
 * reference's default, "top") or GPF_SLIP_BOTH ("both"); dqx, dqy (3, n): in-plane gradients of (rho, jx, jy), NULL =
 * zero.  The time step never takes this route (WallStress / BulkStress call slip = "top" with zero gradients:
 * stress.py:328-344, 452-459 -- gpf_stress_* above, bit-exact); it exists for the reference's analytic tests
 * (tests/test_analytic.py:31-123) and for users of the pure functions.  Closed forms derived from the parabolic
 * velocity profile with Navier slip and the Newtonian stress (tools/derive_slip_closures.py ->
 * csrc/closures_slip_gen.cuh): equal to the reference's expanded polynomials to rounding (<= 1e-12 of the largest
 * component), not bit for bit. */
enum { GPF_SLIP_TOP = 0, GPF_SLIP_BOTH = 1 };
int gpf_stress_ex(const gpf_fv_cfg* cfg, int which, int slip, const double* q, const double* h, const double* Ls,
                  const double* dqx, const double* dqy, double* out, size_t n, void* stream);
/* the same routine evaluated on the HOST (host arrays; no device call): CPU tests of the closed forms */
int gpf_selftest_stress_ex_host(const gpf_fv_cfg* cfg, int which, int slip, const double* q, const double* h,
                                const double* Ls, const double* dqx, const double* dqy, double* out, size_t n);

/* The same two exchanges over peer memory instead of a communication library: every rank owns a
 * small buffer that its peers map through CUDA IPC (one node, NVLink / NVSwitch).  The sending
 * kernel stores the edge rows (the folded record) straight into the receivers' buffers and then
 * publishes a sequence number; the receiving kernel waits for it (bounded, gpf_comm_set_timeout:
 * a dead peer raises the error flag and stops every rank instead of hanging the GPU) and fills the ghost rows (folds the records and
 * advances the loop state).  No reference counterpart (problem.py:676-707 is a local copy).
 *   create on every rank -> all-gather the GPF_IPC_HANDLE_BYTES handles by any means -> connect ->
 *   barrier; then, in the same order on every rank,
 *     halo_send(rel, lo, hi); halo_recv(rel)              lo / hi: neighbour ranks, -1 = none
 *     record_send(kind); finish_recv(mode, recs_out)      kind, mode as reduce_local / finish;
 *                                                         recs_out: device array [world][5] or NULL
 * Exchange and fold counters live on the device, so no call carries a sequence number and
 * gpf_fv_slab_step replays one captured step (passes, both exchanges, fold: 10 kernels) as a CUDA graph. */
#define GPF_IPC_HANDLE_BYTES 64
typedef struct gpf_comm gpf_comm;
int gpf_comm_create(int world, int rank, int Ny, unsigned char* handle_out, gpf_comm** out);
int gpf_comm_connect(gpf_comm* c, const unsigned char* handles /* [world][GPF_IPC_HANDLE_BYTES] */);
int gpf_comm_destroy(gpf_comm* c);
int gpf_comm_error(gpf_comm* c, int* err);       /* synchronises; 1 = a wait timed out */
/* Bound of every device-side wait on a peer (default 30 s, or GPF_COMM_TIMEOUT_S at creation).  A wait that
 * exceeds it raises the error flag; the flag travels to every rank with the next record fold, every rank
 * stops (gpf_fv_scalars.stop) and gpf_fv_get_scalars returns GPF_E_STATE, so stale halo rows never reach
 * a result silently.  Choose it longer than the largest host-side skew between ranks (output, training). */
int gpf_comm_set_timeout(gpf_comm* c, double seconds);
int gpf_fv_halo_send(gpf_fv* s, gpf_comm* c, int rel, int lo, int hi, void* stream);
int gpf_fv_halo_recv(gpf_fv* s, gpf_comm* c, int rel, void* stream);
int gpf_fv_record_send(gpf_fv* s, gpf_comm* c, int kind, void* stream);
int gpf_fv_finish_recv(gpf_fv* s, gpf_comm* c, int mode, double* recs_out, void* stream);
/* n x Problem.update() of this rank's slab (problem.py:509-569) */
int gpf_fv_slab_step(gpf_fv* s, gpf_comm* c, int lo, int hi, int n, double* recs_out, void* stream);

/* Fused exchange: with the neighbours' q buffers mapped (gpf_comm_attach), the pass kernels do the exchange
 * themselves.  The boundary tiles store the cells of row 1 / row Nx (and their y-ghost images) straight into the
 * lower / upper neighbour's ghost row of the same rotating buffer; the general launch of the pass (the second of
 * its two launches, which otherwise returns at once) raises the neighbours' flags and, in the corrector, folds this
 * rank's records and stores them into every rank's table; the thread that issues the TMA loads of a first / last
 * tile row waits for the neighbour's flag of the previous pass.  gpf_fv_slab_step then replays 5 kernels per step
 * (2 + 2 + the fold), none of them an exchange kernel.  On a periodic ring the ranks at the seam count their edge row
 * twice in the kinetic-energy sum (it is also the global ghost row of the reference's padded array, which lives in
 * the neighbour's memory) instead of reducing the received row with one more kernel.
 *   every rank: gpf_fv_peer_export(s, blob) -> all-gather the GPF_PEER_BLOB_BYTES blobs -> gpf_comm_attach(c, s, blobs, lo, hi)
 * The workspace must come from cudaMalloc (a torch tensor of the default caching allocator does): legacy CUDA IPC
 * cannot export stream-ordered or virtual-memory allocations; gpf_fv_peer_export then fails with GPF_E_CUDA and
 * the caller keeps the exchange kernels above.  GPF_SLAB_FUSED=0 keeps them too.
 * gpf_comm_connect_local / gpf_comm_attach_local: the same wiring between communicators and solvers of ONE process
 * (all ranks' slabs on one GPU, driven in lock step pass by pass with gpf_fv_slab_pass: what the single-GPU tests run). */
#define GPF_PEER_BLOB_BYTES 96
int gpf_fv_peer_export(gpf_fv* s, unsigned char* blob);
int gpf_comm_attach(gpf_comm* c, gpf_fv* s, const unsigned char* blobs /* [world][GPF_PEER_BLOB_BYTES] */, int lo, int hi);
/* undo gpf_comm_attach (every rank, or none: the ranks' step graphs must match): gpf_fv_slab_step runs the exchange
 * kernels again.  For a caller whose attach failed on SOME rank. */
int gpf_comm_detach(gpf_comm* c);
int gpf_comm_connect_local(gpf_comm* c, gpf_comm* const* all /* [world] */);
int gpf_comm_attach_local(gpf_comm* c, gpf_fv* s, gpf_fv* lo_solver, int lo, gpf_fv* hi_solver, int hi);
/* one pass (0 predictor, 1 corrector) with the exchange fused in; the step ends with gpf_fv_finish_recv */
int gpf_fv_slab_pass(gpf_fv* s, gpf_comm* c, int corrector, void* stream);

/* --- GP surrogate inference (GaPFlow/models/gp.py) -------------------------------------
 * One handle per GP model (pressure: n_out = 1; wall shear: n_out = 2 outputs sharing one
 * kernel, gp.py:546-552).  Kernel: exp(log_amp) * Matern32(|| exp(-log_scale) . (x - x') ||)
 * (gp.py:576-603).  Inputs are the reference's NORMALISED arrays: Xtrain = database.Xtrain[:,
 * active_dims], Y = Ytrain / Yscale, Xtest = (_Xtest / X_scale)[:, active_dims]; outputs are in
 * normalised units (the caller multiplies by Yscale, Yscale^2 as gp.py:362, 384-386 do). */
typedef struct gpf_gp gpf_gp;
int gpf_gp_workspace_bytes(int M_max, int D, int n_out, size_t* bytes);
int gpf_gp_create(int M_max, int D, int n_out, void* workspace, size_t bytes, gpf_gp** out);
int gpf_gp_destroy(gpf_gp* g);
/* Arithmetic of the O(M)-per-query mean kernel (gpf_gp_mean; the mean part of gpf_gp_mean_var).  GPF_ARITH_FAST (the
 * default; GPF_GP_ARITH=exact in the environment changes the default): exp(-r) by a 32-entry table and a degree-5
 * polynomial, sqrt by one correction of the hardware reciprocal square root -- <= 1e-12 relative on the mean, half the
 * FP64 instructions per kernel evaluation.  GPF_ARITH_EXACT: correctly rounded exp and sqrt.  There is no
 * bit-exact reference for this path (tinygp / JAX's XLA kernels round differently again): its tolerance is 1e-9. */
int gpf_gp_set_arith(gpf_gp* g, int arith);
/* build_gp + the solves of gp.condition (gp.py:268-288, 546-568): K + yerr^2 I, its Cholesky
 * factor, alpha = K^-1 Y.  Xtrain (M, D), Y (M, n_out), log_scale (D): HOST arrays (they come
 * from the host-side database and optimiser).  This is the refactorisation that follows every
 * added MD point (gp.py:290-335).  Synchronises; GPF_E_INVALID if not positive definite. */
int gpf_gp_factor(gpf_gp* g, const double* Xtrain, int M, double log_amp, const double* log_scale,
                  double yerr, const double* Y, void* stream);
/* copies of L (M, M; lower) and alpha (M, n_out) to HOST arrays (tests, gp_*.txt); NULL skips */
int gpf_gp_get_factor(gpf_gp* g, double* L_out, double* alpha_out, void* stream);
/* The training objective of _train (gp.py:290-335) at the parameters of the last gpf_gp_factor:
 *   loss = -sum_outputs log N(Y_o; 0, K + yerr^2 I)   (loss_so / loss_mo, gp.py:306-318)
 * and its gradient grad[0] = d loss / d log_amp, grad[1 + d] = d loss / d log_scale[d] (what
 * jaxopt.ScipyMinimize obtains by autodiff), so that a host optimiser can drive factor + nlml.
 * `work`: device scratch of gpf_gp_train_workspace_bytes(M) bytes (2 M^2 doubles).  Synchronises. */
int gpf_gp_train_workspace_bytes(int M, size_t* bytes);
int gpf_gp_nlml(gpf_gp* g, void* work, size_t bytes, double* loss, double* grad, void* stream);
/* _repredict_mean (gp.py:525-535): Xtest (N, D) device -> mean (n_out, N) device.  If
 * dmean_dx0_max != NULL also max_i d mean_0 / d x_0 (host scalar; Pressure.v_sound's
 * vmap(grad(eos))(Xtest)[:, 0].max(), stress.py:522-537) and the call synchronises. */
int gpf_gp_mean(gpf_gp* g, const double* Xtest, int64_t N, double* mean, double* dmean_dx0_max, void* stream);
/* _repredict_mean_var (gp.py:509-522): also var (N) = amp + noise - sum (L^-1 Ks)^2.  noise: device
 * array (N) or NULL for the scalar `noise_scalar` (tinygp's conditioned-GP jitter: an input, not
 * guessed).  var_max / arg_max (host; gp.py:406, 424: first maximum, ghost cells included) may be
 * NULL; if either is given the call synchronises. */
int gpf_gp_mean_var(gpf_gp* g, const double* Xtest, int64_t N, const double* noise, double noise_scalar,
                    double* mean, double* var, double* var_max, int64_t* arg_max, void* stream);
/* Feature matrix of the current solution straight from the FV solver's storage (gp.py:223-232,
 * stress.py:194-197): X (Np, D) device, Np = (Nx+2)(Ny+2), row i*(Ny+2)+j; column k is feature
 * active_dims[k] of [rho, jx, jy, h, dh/dx, dh/dy, extra] divided by x_scale[active_dims[k]].
 * x_scale (7) and active_dims (D) are host arrays. */
int gpf_gp_features(gpf_fv* s, const double* x_scale, const int32_t* active_dims, int D, double* X, void* stream);
/* the same of buffer `rel`: 0 the current q, 1 the predictor's output (the corrector pass's input), 2 the
 * corrector's output before gpf_fv_finish made it current (q = (q + q0)/2, ghosts filled: the field
 * _post_update evaluates v_sound on, problem.py:571-586) */
int gpf_gp_features_of(gpf_fv* s, int rel, const double* x_scale, const int32_t* active_dims, int D, double* X,
                       void* stream);

/* --- instrumentation ------------------------------------------------------------- */
/* number of kernels this library has launched since load (bench.py's gpu_launches) */
int64_t gpf_launch_count(void);
/* average device time (ms) of the last gpf_fv_step() batch's predictor / corrector
 * kernels measured with CUDA events on the launching stream; 0 if timing disabled */
int gpf_fv_set_timing(gpf_fv* s, int enable);
int gpf_fv_get_timing(gpf_fv* s, double* ms_predictor, double* ms_corrector, int64_t* launches);

/* Self-test: n seeded operand pairs (structured mantissas, |x| in [2^-90, 2^90]); counts where
 * the kernels' branch-free reciprocal / quotient differ from the IEEE-754 operations
 * (__drcp_rn, __ddiv_rn).  mismatches[0]: reciprocal, [1]: quotient.  Synchronises. */
int gpf_selftest_division(uint64_t n, uint64_t seed, uint64_t mismatches[2], void* stream);

/* Self-test, HOST arithmetic only (no device call): the O(log M) mean of a one-input, one-output model
 * (csrc/gp_mean1d.h: sorted training points, blockwise running sums, bisection per query) evaluated by
 * the same inline routine the device kernel calls.  gpf_gp_mean takes this route whenever D = 1 and
 * n_out = 1 (GPF_GP_MEAN1D=0 restores the O(M) kernel).  Xtrain (M), alpha (M), Xtest (N): normalised,
 * unscaled inputs as for gpf_gp_factor / gpf_gp_mean; mean (N), dmean_dx0 (N) or NULL. */
int gpf_selftest_mean1d_host(const double* Xtrain, int M, double log_amp, double log_scale, const double* alpha,
                             const double* Xtest, int64_t N, double* mean, double* dmean_dx0);

/* Self-test, HOST arithmetic only: the O(log M) predictive variance of a one-input model (csrc/gp_var1d.h:
 * the one-dimensional Matern-3/2 process as a two-state Markov process, Kalman sweeps from both ends of the
 * sorted training points, two-filter combination per query), by the same inline routine as the device kernel
 * that gpf_gp_mean_var uses for D = 1 (GPF_GP_VAR1D=0 restores the dense solve).
 * var (N) = posterior variance + noise, i.e. gp.py:509-522's `Kss - sum(v**2)`. */
int gpf_selftest_var1d_host(const double* Xtrain, int M, double log_amp, double log_scale, double yerr,
                            const double* Xtest, int64_t N, double noise, double* var);

#ifdef __cplusplus
}
#endif
#endif /* GAPFLOW_B200_H */
