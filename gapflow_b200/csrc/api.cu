// extern "C" boundary of libgapflow_b200.so (see include/gapflow_b200.h).
#include <cmath>
#include <cstring>
#include <new>
#include <vector>
#include <string>
#include "fv_host.h"
#include "gp_host.h"
#include "slab_comm.cuh"

namespace gpf {
static thread_local std::string t_err;
void set_error(const std::string& m) { t_err = m; }
int cuda_fail(cudaError_t e, const char* what) {
    t_err = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
    return GPF_E_CUDA;
}
}  // namespace gpf

using namespace gpf;

struct gpf_fv {
    FvSolver S;
};
struct gpf_gp {
    GpModel G;
};
struct gpf_comm {
    SlabComm C;
};

#define GPF_TRY try {
#define GPF_CATCH                                                            \
    }                                                                        \
    catch (const std::bad_alloc&) { set_error("out of host memory"); return GPF_E_NOMEM; } \
    catch (const std::exception& e) { set_error(e.what()); return GPF_E_INVALID; }        \
    catch (...) { set_error("unknown C++ exception"); return GPF_E_INVALID; }

extern "C" {

const char* gpf_last_error(void) { return t_err.c_str(); }
int gpf_version(void) { return 100; }
int64_t gpf_launch_count(void) { return (int64_t)g_launches.load(); }

// masks[s][k][c]: s: 0 xE, 1 xW, 2 yS, 3 yN; k: 0 P, 1 D, 2 N.
// Reference (problem.py:681-707, 736-754):
//   ghost x0 (index 0):    gate all(xE_P); assigned through masks xE_{D,N}; values from
//                           comps selected by xW_{D,N}, target xW_D_val, adjacent row 1
//   ghost x1 (index -1):   gate all(xW_P); assigned through xW_*; values from xE_*, xE_D_val, row -2
//   ghost y0 (index 0):    gate all(yS_P); assigned through yS_*; values from yN_*, yN_D_val, col 1
//   ghost y1 (index -1):   gate all(yN_P); assigned through yN_*; values from yS_*, yS_D_val, col -2
// NumPy assigns value array (k_val, 1, n) into (k_dst, 1, n): legal iff k_val == k_dst
// or k_val == 1 (broadcast).
int gpf_fv_cfg_set_bc(gpf_fv_cfg* cfg, const uint8_t masks[4][3][3], const double d_val[4]) {
    GPF_TRY
    if (!cfg || !masks || !d_val) { set_error("null argument"); return GPF_E_INVALID; }
    // side -> (assign masks index, value masks index)
    const int assign[4] = {0, 1, 2, 3};   // x0: xE, x1: xW, y0: yS, y1: yN
    const int value[4] = {1, 0, 3, 2};    // x0: xW, x1: xE, y0: yN, y1: yS
    for (int s = 0; s < 4; ++s) {
        const int A = assign[s], Vv = value[s];
        cfg->bc_val[s] = d_val[Vv];
        bool allP = masks[A][0][0] && masks[A][0][1] && masks[A][0][2];
        for (int c = 0; c < 3; ++c) { cfg->bc_type[s][c] = GPF_BC_P; cfg->bc_src[s][c] = c; }
        if (allP) continue;
        bool assigned[3] = {false, false, false};
        for (int k = 1; k <= 2; ++k) {       // D then N, as the reference assigns them
            int dst[3], nd = 0, src[3], ns = 0;
            for (int c = 0; c < 3; ++c) { if (masks[A][k][c]) dst[nd++] = c; if (masks[Vv][k][c]) src[ns++] = c; }
            if (nd == 0 && ns <= 1) continue; // empty selection on the left: a (0|1, 1, n) value broadcasts into it, a no-op
            if (!(ns == nd || ns == 1)) {     // includes nd == 0 with ns >= 2: NumPy cannot broadcast (2, 1, n) into (0, 1, n)
                set_error("boundary masks: the reference's ghost-cell assignment would raise (shape mismatch between E/W or S/N masks)");
                return GPF_E_INVALID;
            }
            for (int t = 0; t < nd; ++t) {
                cfg->bc_type[s][dst[t]] = (k == 1) ? GPF_BC_D : GPF_BC_N;
                cfg->bc_src[s][dst[t]] = (ns == 1) ? src[0] : src[t];
                assigned[dst[t]] = true;
            }
        }
        for (int c = 0; c < 3; ++c)
            if (!assigned[c]) {
                set_error("boundary masks: a side mixes periodic with D/N components; the reference leaves those ghost cells undefined");
                return GPF_E_INVALID;
            }
    }
    return GPF_OK;
    GPF_CATCH
}

int gpf_fv_workspace_bytes(const gpf_fv_cfg* cfg, size_t* bytes) {
    GPF_TRY
    if (!cfg || !bytes) { set_error("null argument"); return GPF_E_INVALID; }
    FvParams P;
    int rc = pack_params(*cfg, &P);
    if (rc) return rc;
    *bytes = FvSolver::workspace_bytes(*cfg);
    return GPF_OK;
    GPF_CATCH
}

int gpf_fv_create(const gpf_fv_cfg* cfg, void* workspace, size_t bytes, gpf_fv** out) {
    GPF_TRY
    if (!cfg || !workspace || !out) { set_error("null argument"); return GPF_E_INVALID; }
    gpf_fv* h = new gpf_fv();
    int rc = h->S.init(*cfg, workspace, bytes);
    if (rc) { delete h; return rc; }
    *out = h;
    return GPF_OK;
    GPF_CATCH
}

int gpf_fv_destroy(gpf_fv* s) {
    GPF_TRY
    if (!s) return GPF_OK;
    s->S.destroy();
    delete s;
    return GPF_OK;
    GPF_CATCH
}

#define NEED(s) if (!(s)) { set_error("null handle"); return GPF_E_INVALID; }

int gpf_fv_upload(gpf_fv* s, int field, const double* data, int ncomp, int on_device, void* stream) {
    GPF_TRY NEED(s && data) return s->S.upload(field, data, ncomp, on_device, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_download(gpf_fv* s, int field, double* data, int ncomp, int on_device, void* stream) {
    GPF_TRY NEED(s && data) return s->S.download(field, data, ncomp, on_device, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_pre_run(gpf_fv* s, int reset_ekin_old, void* stream) {
    GPF_TRY NEED(s) return s->S.pre_run(reset_ekin_old, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_step(gpf_fv* s, int n, int until_converged, void* stream) {
    GPF_TRY NEED(s) return s->S.step(n, until_converged, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_fill_ghosts(gpf_fv* s, void* stream) {
    GPF_TRY NEED(s) return s->S.fill_ghosts((cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_get_scalars(gpf_fv* s, gpf_fv_scalars* out, void* stream) {
    GPF_TRY NEED(s && out) return s->S.get_scalars(out, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_set_scalars(gpf_fv* s, const gpf_fv_scalars* in, void* stream) {
    GPF_TRY NEED(s && in) return s->S.set_scalars(in, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_mass(gpf_fv* s, double* mass, void* stream) {
    GPF_TRY NEED(s && mass) return s->S.mass(mass, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_closures(gpf_fv* s, double* pressure, double* wall_xz, double* wall_yz, double* bulk, int on_device,
                    void* stream) {
    GPF_TRY NEED(s) return s->S.closures(pressure, wall_xz, wall_yz, bulk, on_device, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_closures_of(gpf_fv* s, int which, double* pressure, double* wall_xz, double* wall_yz, double* bulk,
                       int on_device, void* stream) {
    GPF_TRY NEED(s)
    if (which != 0 && which != 1) { set_error("which must be 0 (current q) or 1 (corrector-pass input)"); return GPF_E_INVALID; }
    return s->S.closures(pressure, wall_xz, wall_yz, bulk, on_device, (cudaStream_t)stream, which ? 2 : 0);
    GPF_CATCH
}
int gpf_fv_set_viscosity(gpf_fv* s, int piezo, const double* piezo_par, int thinning, const double* thin_par,
                         const double* x, const double* y) {
    GPF_TRY NEED(s) return s->S.set_viscosity(piezo, piezo_par, thinning, thin_par, x, y); GPF_CATCH
}
int gpf_fv_viscosity(gpf_fv* s, int which, double* eta, int on_device, void* stream) {
    GPF_TRY NEED(s && eta)
    if (which != 0 && which != 1) { set_error("which must be 0 (current q) or 1 (corrector-pass input)"); return GPF_E_INVALID; }
    return s->S.viscosity(which, eta, on_device, (cudaStream_t)stream);
    GPF_CATCH
}
int gpf_fv_update_host(gpf_fv* s, double* q_host, int n, gpf_fv_scalars* out, void* stream) {
    GPF_TRY
    NEED(s && q_host)
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    if (n == 1) {
        // copies overlapped with the passes chunk by chunk; GPF_E_STATE: configuration not eligible
        rc = s->S.update_host_pipelined(q_host, st);
        if (rc == GPF_OK) {
            if (out) return s->S.get_scalars(out, st);
            GPF_CUDA(cudaStreamSynchronize(st));
            return GPF_OK;
        }
        if (rc != GPF_E_STATE) return rc;
    }
    if ((rc = s->S.upload(GPF_F_Q, q_host, 3, 0, st))) return rc;
    s->S.dirty_scalars = false;    // loop scalars (dt, ekin_old) are kept; only q is replaced
    if ((rc = s->S.step(n, 0, st))) return rc;
    if ((rc = s->S.download(GPF_F_Q, q_host, 3, 0, st))) return rc;
    if (out) return s->S.get_scalars(out, st);
    GPF_CUDA(cudaStreamSynchronize(st));
    return GPF_OK;
    GPF_CATCH
}
int gpf_fv_pass(gpf_fv* s, int corrector, void* stream) {
    GPF_TRY NEED(s) return s->S.pass(corrector, GPF_PART_ALL, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_pass_part(gpf_fv* s, int corrector, int part, void* stream) {
    GPF_TRY NEED(s) return s->S.pass(corrector, part, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_set_run_mode(gpf_fv* s, int until_converged) {
    GPF_TRY NEED(s)
    const int uc = until_converged ? 1 : 0;
    if (s->S.run_mode_uc != uc) { s->S.run_mode_uc = uc; s->S.drop_graphs(); }
    return GPF_OK;
    GPF_CATCH
}
int gpf_fv_halo_pack(gpf_fv* s, int rel, double* sendbuf, void* stream) {
    GPF_TRY NEED(s && sendbuf) return s->S.halo_pack(rel, sendbuf, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_halo_unpack(gpf_fv* s, int rel, const double* recvbuf, void* stream) {
    GPF_TRY NEED(s && recvbuf) return s->S.halo_unpack(rel, recvbuf, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_reduce_local(gpf_fv* s, int kind, double* rec5, void* stream) {
    GPF_TRY NEED(s && rec5) return s->S.reduce_local(kind, rec5, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_finish(gpf_fv* s, int mode, const double* recs, int nrec, void* stream) {
    GPF_TRY NEED(s) return s->S.finish(mode, recs, nrec, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_set_surrogate(gpf_fv* s, const double* pressure, double pressure_scale, const double* shear_xz,
                         double shear_xz_scale, const double* shear_yz, double shear_yz_scale) {
    GPF_TRY NEED(s) return s->S.set_surrogate(pressure, pressure_scale, shear_xz, shear_xz_scale, shear_yz, shear_yz_scale); GPF_CATCH
}
int gpf_fv_finish_surrogate(gpf_fv* s, int mode, double* recs, int nrec, double c2, void* stream) {
    GPF_TRY NEED(s) return s->S.finish_surrogate(mode, recs, nrec, c2, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_set_timing(gpf_fv* s, int enable) {
    GPF_TRY NEED(s) s->S.timing = enable != 0; s->S.ev_steps = 0; return GPF_OK; GPF_CATCH
}
int gpf_fv_get_timing(gpf_fv* s, double* ms_predictor, double* ms_corrector, int64_t* launches) {
    GPF_TRY NEED(s)
    long long l = 0;
    int rc = s->S.get_timing(ms_predictor, ms_corrector, &l);
    if (launches) *launches = l;
    return rc;
    GPF_CATCH
}

int gpf_eos_pressure(const gpf_fv_cfg* cfg, const double* dens, double* out, size_t n, void* stream) {
    GPF_TRY NEED(cfg && (n == 0 || (dens && out))) return launch_eos(*cfg, dens, out, n, false, (cudaStream_t)stream); GPF_CATCH
}
int gpf_eos_sound_velocity(const gpf_fv_cfg* cfg, const double* dens, double* out, size_t n, void* stream) {
    GPF_TRY NEED(cfg && (n == 0 || (dens && out))) return launch_eos(*cfg, dens, out, n, true, (cudaStream_t)stream); GPF_CATCH
}
int gpf_stress_bottom(const gpf_fv_cfg* cfg, const double* q, const double* h, const double* Ls, double* out,
                      size_t n, void* stream) {
    GPF_TRY NEED(cfg && (n == 0 || (q && h && out))) return launch_stress(*cfg, 0, q, h, Ls, out, n, (cudaStream_t)stream); GPF_CATCH
}
int gpf_stress_top(const gpf_fv_cfg* cfg, const double* q, const double* h, const double* Ls, double* out, size_t n,
                   void* stream) {
    GPF_TRY NEED(cfg && (n == 0 || (q && h && out))) return launch_stress(*cfg, 1, q, h, Ls, out, n, (cudaStream_t)stream); GPF_CATCH
}
int gpf_stress_avg(const gpf_fv_cfg* cfg, const double* q, const double* h, const double* Ls, double* out, size_t n,
                   void* stream) {
    GPF_TRY NEED(cfg && (n == 0 || (q && h && out))) return launch_stress(*cfg, 2, q, h, Ls, out, n, (cudaStream_t)stream); GPF_CATCH
}

int gpf_stress_ex(const gpf_fv_cfg* cfg, int which, int slip, const double* q, const double* h, const double* Ls,
                  const double* dqx, const double* dqy, double* out, size_t n, void* stream) {
    GPF_TRY NEED(cfg && (n == 0 || (q && h && out)))
    if (slip != GPF_SLIP_TOP && slip != GPF_SLIP_BOTH) { set_error("slip: GPF_SLIP_TOP or GPF_SLIP_BOTH"); return GPF_E_INVALID; }
    return launch_stress_ex(*cfg, which, slip == GPF_SLIP_BOTH, q, h, Ls, dqx, dqy, out, n, (cudaStream_t)stream);
    GPF_CATCH
}
int gpf_selftest_stress_ex_host(const gpf_fv_cfg* cfg, int which, int slip, const double* q, const double* h, const double* Ls,
                                const double* dqx, const double* dqy, double* out, size_t n) {
    GPF_TRY NEED(cfg && (n == 0 || (q && h && out)))
    if (slip != GPF_SLIP_TOP && slip != GPF_SLIP_BOTH) { set_error("slip: GPF_SLIP_TOP or GPF_SLIP_BOTH"); return GPF_E_INVALID; }
    return stress_ex_host(*cfg, which, slip == GPF_SLIP_BOTH, q, h, Ls, dqx, dqy, out, n);
    GPF_CATCH
}

int gpf_gp_workspace_bytes(int M_max, int D, int n_out, size_t* bytes) {
    GPF_TRY
    if (!bytes || M_max < 1 || D < 1 || D > 8 || n_out < 1 || n_out > 2) { set_error("bad GP dimensions (D <= 8, n_out <= 2)"); return GPF_E_INVALID; }
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaGetLastError();
    if (sms <= 0) sms = 148;
    *bytes = GpModel::workspace_bytes(M_max, D, n_out, GpModel::var_slots(sms));
    return GPF_OK;
    GPF_CATCH
}
int gpf_gp_create(int M_max, int D, int n_out, void* workspace, size_t bytes, gpf_gp** out) {
    GPF_TRY
    if (!workspace || !out) { set_error("null argument"); return GPF_E_INVALID; }
    gpf_gp* h = new gpf_gp();
    int rc = h->G.init(M_max, D, n_out, workspace, bytes);
    if (rc) { delete h; return rc; }
    *out = h;
    return GPF_OK;
    GPF_CATCH
}
int gpf_gp_set_arith(gpf_gp* g, int arith) {
    GPF_TRY NEED(g)
    if (arith != GPF_ARITH_EXACT && arith != GPF_ARITH_FAST) { set_error("arith must be GPF_ARITH_EXACT or GPF_ARITH_FAST"); return GPF_E_INVALID; }
    g->G.fast_arith = arith == GPF_ARITH_FAST;
    return GPF_OK;
    GPF_CATCH
}
int gpf_gp_destroy(gpf_gp* g) {
    GPF_TRY
    if (g) {
        if (g->G.side) { cudaStreamSynchronize(g->G.side); cudaStreamDestroy(g->G.side); }
        if (g->G.ev_strip) cudaEventDestroy(g->G.ev_strip);
        if (g->G.ev_rest) cudaEventDestroy(g->G.ev_rest);
    }
    delete g;
    return GPF_OK;
    GPF_CATCH
}
int gpf_gp_factor(gpf_gp* g, const double* Xtrain, int M, double log_amp, const double* log_scale, double yerr,
                  const double* Y, void* stream) {
    GPF_TRY NEED(g && Xtrain && log_scale && Y) return g->G.factor(Xtrain, M, log_amp, log_scale, yerr, Y, (cudaStream_t)stream); GPF_CATCH
}
int gpf_gp_get_factor(gpf_gp* g, double* L_out, double* alpha_out, void* stream) {
    GPF_TRY NEED(g) return g->G.get_factor(L_out, alpha_out, (cudaStream_t)stream); GPF_CATCH
}
int gpf_comm_create(int world, int rank, int Ny, unsigned char* handle_out, gpf_comm** out) {
    GPF_TRY
    NEED(handle_out && out)
    static_assert(sizeof(cudaIpcMemHandle_t) == GPF_IPC_HANDLE_BYTES, "IPC handle size");
    gpf_comm* c = new gpf_comm();
    int rc = c->C.create(world, rank, Ny);
    if (rc) { delete c; return rc; }
    memcpy(handle_out, &c->C.handle, GPF_IPC_HANDLE_BYTES);
    *out = c;
    return GPF_OK;
    GPF_CATCH
}
int gpf_comm_connect(gpf_comm* c, const unsigned char* handles) {
    GPF_TRY NEED(c && handles) return c->C.connect(handles); GPF_CATCH
}
int gpf_comm_destroy(gpf_comm* c) {
    GPF_TRY NEED(c) c->C.destroy(); delete c; return GPF_OK; GPF_CATCH
}
int gpf_comm_error(gpf_comm* c, int* err) {
    GPF_TRY
    NEED(c && err)
    if (cudaMemcpy(err, c->C.err(), sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) { set_error("cudaMemcpy failed in gpf_comm_error"); return GPF_E_CUDA; }
    return GPF_OK;
    GPF_CATCH
}
int gpf_fv_peer_export(gpf_fv* s, unsigned char* blob) {
    GPF_TRY NEED(s && blob) return s->S.peer_export(blob); GPF_CATCH
}
int gpf_comm_attach(gpf_comm* c, gpf_fv* s, const unsigned char* blobs, int lo, int hi) {
    GPF_TRY NEED(c && s && blobs) return c->C.attach(s->S, blobs, lo, hi); GPF_CATCH
}
int gpf_comm_detach(gpf_comm* c) {
    GPF_TRY NEED(c) return c->C.detach(); GPF_CATCH
}
int gpf_comm_connect_local(gpf_comm* c, gpf_comm* const* all) {
    GPF_TRY NEED(c && all)
    std::vector<SlabComm*> v(c->C.lay.world);
    for (int r = 0; r < c->C.lay.world; ++r) { NEED(all[r]) v[r] = &all[r]->C; }
    return c->C.connect_local(v.data());
    GPF_CATCH
}
int gpf_comm_attach_local(gpf_comm* c, gpf_fv* s, gpf_fv* lo_solver, int lo, gpf_fv* hi_solver, int hi) {
    GPF_TRY NEED(c && s) return c->C.attach_local(s->S, lo_solver ? &lo_solver->S : nullptr, lo, hi_solver ? &hi_solver->S : nullptr, hi); GPF_CATCH
}
int gpf_fv_slab_pass(gpf_fv* s, gpf_comm* c, int corrector, void* stream) {
    GPF_TRY NEED(s && c) return s->S.slab_pass(c->C, corrector, (cudaStream_t)stream); GPF_CATCH
}
int gpf_comm_set_timeout(gpf_comm* c, double seconds) {
    GPF_TRY NEED(c) return c->C.set_timeout(seconds); GPF_CATCH
}
int gpf_fv_halo_send(gpf_fv* s, gpf_comm* c, int rel, int lo, int hi, void* stream) {
    GPF_TRY NEED(s && c) return s->S.halo_send(c->C, rel, lo, hi, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_halo_recv(gpf_fv* s, gpf_comm* c, int rel, void* stream) {
    GPF_TRY NEED(s && c) return s->S.halo_recv(c->C, rel, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_record_send(gpf_fv* s, gpf_comm* c, int kind, void* stream) {
    GPF_TRY NEED(s && c) return s->S.record_send(c->C, kind, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_finish_recv(gpf_fv* s, gpf_comm* c, int mode, double* recs_out, void* stream) {
    GPF_TRY NEED(s && c) return s->S.finish_recv(c->C, mode, recs_out, (cudaStream_t)stream); GPF_CATCH
}
int gpf_fv_slab_step(gpf_fv* s, gpf_comm* c, int lo, int hi, int n, double* recs_out, void* stream) {
    GPF_TRY NEED(s && c) return s->S.slab_step(c->C, lo, hi, n, recs_out, (cudaStream_t)stream); GPF_CATCH
}
int gpf_gp_train_workspace_bytes(int M, size_t* bytes) {
    GPF_TRY NEED(bytes && M > 0) *bytes = gpf::GpModel::train_workspace_bytes(M); return GPF_OK; GPF_CATCH
}
int gpf_gp_nlml(gpf_gp* g, void* work, size_t bytes, double* loss, double* grad, void* stream) {
    GPF_TRY NEED(g && work) return g->G.nlml(work, bytes, loss, grad, (cudaStream_t)stream); GPF_CATCH
}
int gpf_gp_mean(gpf_gp* g, const double* Xtest, int64_t N, double* mean, double* dmean_dx0_max, void* stream) {
    GPF_TRY NEED(g && (N == 0 || (Xtest && mean))) return g->G.mean(Xtest, N, mean, dmean_dx0_max, (cudaStream_t)stream); GPF_CATCH
}
int gpf_gp_mean_var(gpf_gp* g, const double* Xtest, int64_t N, const double* noise, double noise_scalar, double* mean,
                    double* var, double* var_max, int64_t* arg_max, void* stream) {
    GPF_TRY NEED(g && (N == 0 || (Xtest && mean && var)))
    long long im = 0;
    int rc = g->G.mean_var(Xtest, N, noise, noise_scalar, mean, var, var_max, arg_max ? &im : nullptr, (cudaStream_t)stream);
    if (arg_max) *arg_max = im;
    return rc;
    GPF_CATCH
}
int gpf_selftest_mean1d_host(const double* Xtrain, int M, double log_amp, double log_scale, const double* alpha,
                             const double* Xtest, int64_t N, double* mean, double* dmean_dx0) {
    GPF_TRY
    if (!Xtrain || !alpha || !Xtest || !mean || M < 1 || N < 0) { set_error("bad arguments"); return GPF_E_INVALID; }
    const double s = exp(-log_scale), amp = exp(log_amp);
    std::vector<double> u(M);
    for (int j = 0; j < M; ++j) u[j] = Xtrain[j] * s;
    Mean1DHost H;
    H.build(u.data(), alpha, M);
    const Mean1D T = H.view();
    for (int64_t i = 0; i < N; ++i) {
        double m, g;
        mean1d_eval(T, Xtest[i] * s, &m, &g);
        mean[i] = amp * m;
        if (dmean_dx0) dmean_dx0[i] = -3.0 * amp * s * g;
    }
    return GPF_OK;
    GPF_CATCH
}
int gpf_selftest_var1d_host(const double* Xtrain, int M, double log_amp, double log_scale, double yerr,
                            const double* Xtest, int64_t N, double noise, double* var) {
    GPF_TRY
    if (!Xtrain || !Xtest || !var || M < 1 || N < 0) { set_error("bad arguments"); return GPF_E_INVALID; }
    const double s = exp(-log_scale);
    std::vector<double> u(M);
    for (int j = 0; j < M; ++j) u[j] = Xtrain[j] * s;
    Var1DHost H;
    H.build(u.data(), M, exp(log_amp), yerr);
    const Var1D T = H.view();
    for (int64_t i = 0; i < N; ++i) var[i] = var1d_eval(T, Xtest[i] * s) + noise;
    return GPF_OK;
    GPF_CATCH
}
int gpf_gp_features(gpf_fv* s, const double* x_scale, const int32_t* active_dims, int D, double* X, void* stream) {
    GPF_TRY NEED(s && x_scale && active_dims && X) return launch_gp_features(s->S, x_scale, active_dims, D, X, (cudaStream_t)stream); GPF_CATCH
}
int gpf_gp_features_of(gpf_fv* s, int rel, const double* x_scale, const int32_t* active_dims, int D, double* X, void* stream) {
    GPF_TRY NEED(s && x_scale && active_dims && X) return launch_gp_features(s->S, x_scale, active_dims, D, X, (cudaStream_t)stream, rel); GPF_CATCH
}

int gpf_selftest_division(uint64_t n, uint64_t seed, uint64_t mismatches[2], void* stream) {
    GPF_TRY NEED(mismatches)
    unsigned long long o[2] = {0, 0};
    int rc = launch_selftest_div(n, seed, o, (cudaStream_t)stream);
    mismatches[0] = o[0]; mismatches[1] = o[1];
    return rc;
    GPF_CATCH
}

}  // extern "C"
