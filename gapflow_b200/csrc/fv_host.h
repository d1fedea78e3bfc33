// Host-side solver object behind the opaque gpf_fv handle.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <atomic>
#include <string>
#include <vector>
#include "gpf_types.h"

namespace gpf {

extern std::atomic<long long> g_launches;
void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what);

#define GPF_CUDA(x)                                        \
    do {                                                   \
        cudaError_t _e = (x);                              \
        if (_e != cudaSuccess) return cuda_fail(_e, #x);   \
    } while (0)

// reduction result of the "refresh" path (scalars of the current q)
struct RefreshOut {
    double ekin, vmax, vsound, mass;
    int has_nan, has_neg;
};

struct FvSolver {
    gpf_fv_cfg cfg;
    FvParams P;
    int device;
    // carved from the caller's workspace
    double* qbuf[3];
    double* topo;      // 3 planes
    double* extra;     // 1 plane
    double* eta = nullptr;       // 1 plane: per-cell shear viscosity (cfg.visc_field), rewritten before every pass
    double* gradtab = nullptr;   // np.gradient coefficient tables: 3 (Nx+2) + 3 (Ny+2) doubles
    ViscModel VM;
    bool visc_set = false;
    CUtensorMap eta_map;
    int set_viscosity(int piezo, const double* pp, int thin, const double* tp, const double* x, const double* y);
    int launch_viscosity(int rel, cudaStream_t s);
    int viscosity(int which, double* out, int on_device, cudaStream_t s);
    // GP surrogate closures (gpf_fv_set_surrogate): normalised predictive means of the pass's input, dense
    // over the padded array; null = the fixed-form law
    const double* sg_p = nullptr;    // (Np)
    const double* sg_xz = nullptr;   // (2, Np): lower, upper wall
    const double* sg_yz = nullptr;
    double sg_scale[3] = {1.0, 1.0, 1.0};
    int set_surrogate(const double* p, double ps, const double* xz, double xzs, const double* yz, double yzs);
    int finish_surrogate(int mode, double* recs, int nrec, double c2, cudaStream_t s);
    DevState* st;
    Partial* partial;
    int* pflags;
    RefreshOut* refresh;
    int npartial;      // capacity of partial/pflags
    int* redo;         // [2] device flags, see fv_pass_kernel
    int nrec_corr = 0; // record slots written by the last corrector pass (all its launches)
    cudaGraphExec_t step_graph[2] = {nullptr, nullptr};   // one captured step, [until_converged]
    void drop_graphs();
    CUtensorMap qmap[3], q0map[3], topo_map, topo2_map, ls_map;
    int topo_fast, ls_fast;   // range flags of the uploaded topography / slip length (see FvParams::static_fast)
    bool dirty_scalars;   // q changed from the host since scalars were computed
    bool refresh_valid;   // `refresh` holds the scalars of the current q (all ranks folded)
    double** qtable;      // device copy of qbuf[3]
    // timing
    bool timing;
    std::vector<cudaEvent_t> ev;   // 3 per step
    int ev_steps;

    static size_t workspace_bytes(const gpf_fv_cfg& c);
    int init(const gpf_fv_cfg& c, void* ws, size_t bytes);
    int destroy();
    int upload(int field, const double* data, int ncomp, int on_device, cudaStream_t s);
    int download(int field, double* data, int ncomp, int on_device, cudaStream_t s);
    int pre_run(int reset_ekin_old, cudaStream_t s);
    int step(int n, int until_converged, cudaStream_t s);
    int fill_ghosts(cudaStream_t s);
    int refresh_scalars(cudaStream_t s);
    int get_scalars(gpf_fv_scalars* out, cudaStream_t s);
    int set_scalars(const gpf_fv_scalars* in, cudaStream_t s);
    int mass(double* m, cudaStream_t s);
    int closures(double* p, double* wxz, double* wyz, double* bulk, int on_device, cudaStream_t s, int rel = 0);
    int pass(int corrector, int part, cudaStream_t s);
    int halo_pack(int rel, double* send, cudaStream_t s);
    int halo_unpack(int rel, const double* recv, cudaStream_t s);
    int reduce_local(int kind, double* rec5, cudaStream_t s);
    int halo_send(struct SlabComm& c, int rel, int lo, int hi, cudaStream_t s);
    int halo_recv(struct SlabComm& c, int rel, cudaStream_t s);
    int record_send(struct SlabComm& c, int kind, cudaStream_t s);
    int finish_recv(struct SlabComm& c, int mode, double* recs_out, cudaStream_t s, int until_converged = 0);
    int slab_step_once(struct SlabComm& c, int lo, int hi, double* recs_out, cudaStream_t s);
    int slab_pass(struct SlabComm& c, int corrector, cudaStream_t s);      // one pass with the exchange fused in
    int peer_export(unsigned char* blob);                                  // GPF_PEER_BLOB_BYTES describing the q buffers
    int slab_step(struct SlabComm& c, int lo, int hi, int n, double* recs_out, cudaStream_t s);
    bool graph_stale = false;
    int run_mode_uc = 0;     // gpf_fv_set_run_mode: pass / finish / slab-step launches honour the run-loop condition on the device
    int slab_launches = 0;
    int finish(int mode, const double* recs, int nrec, cudaStream_t s);
    // host-buffer step with the copies overlapped chunk by chunk (gpf_fv_update_host, n == 1)
    cudaStream_t cp_stream[2] = {nullptr, nullptr};
    std::vector<cudaEvent_t> cp_event;
    int update_host_pipelined(double* q_host, cudaStream_t s);
    int get_timing(double* ms_pred, double* ms_corr, long long* launches);
    double* field_ptr(int field, int* ncomp_max);
    int read_cur(cudaStream_t s, int* cur);
};

// kernel argument block of the pass kernels (fv_pass_kernel.cuh)
struct PassArgs {
    CUtensorMap qmap[3];       // box (BY, BX, 3) of each rotating q buffer
    CUtensorMap q0map[3];      // box (TY, TX, 3): own cells only
    CUtensorMap topo_map;      // box (BY, BX, 3)
    CUtensorMap topo2_map;     // box (BY, BX, 2): h, dh/dx
    CUtensorMap ls_map;        // box (BY, BX, 1)
    CUtensorMap eta_map;       // box (BY, BX, 1): per-cell shear viscosity (viscosity-field variants only)
    FvParams P;
    double* qbuf[3];
    DevState* st;
    Partial* partial;
    int* pflags;
    int* redo;                 // [2]: pass (0 predictor, 1 corrector) must be redone by the general kernel
    int until_converged;
    int ntiles;                // tiles of this launch (virtual index t -> tile row (t / nty) * bx_mul + bx_off)
    int bx_mul, bx_off;
    int rec0;                  // first record slot of this launch
    SlabX X;                   // fused slab exchange (X.on = 0: plain pass)
    // surrogate variant only (SG): GP means replacing the pressure / the own shear component of the wall stresses
    const double* sg_p;
    const double* sg_xz;
    const double* sg_yz;
    double sg_scale[3];
};

// parameter packing shared with the pure-function entry points
int pack_params(const gpf_fv_cfg& c, FvParams* P);
Layout make_layout(int Nx, int Ny);

// pure functions
int launch_eos(const gpf_fv_cfg& c, const double* dens, double* out, size_t n, bool sound, cudaStream_t s);
int launch_stress(const gpf_fv_cfg& c, int which, const double* q, const double* h, const double* Ls,
                  double* out, size_t n, cudaStream_t s);

int launch_stress_ex(const gpf_fv_cfg& c, int which, int slip_both, const double* q, const double* h, const double* Ls,
                     const double* dqx, const double* dqy, double* out, size_t n, cudaStream_t s);
int stress_ex_host(const gpf_fv_cfg& c, int which, int slip_both, const double* q, const double* h, const double* Ls,
                   const double* dqx, const double* dqy, double* out, size_t n);

int launch_selftest_div(unsigned long long n, unsigned long long seed, unsigned long long* out_host, cudaStream_t s);

}  // namespace gpf
