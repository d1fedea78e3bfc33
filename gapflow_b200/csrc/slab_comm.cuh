// Slab neighbours and record fold over peer memory (NVLink P2P stores), no NCCL on the step path.
//
// Each rank owns one small cudaMalloc'd buffer that its peers map through CUDA IPC:
//   halo [2 parities][2 sides][3][Ny+2]   edge rows written by the neighbours (side 0: from the lower
//                                          neighbour -> ghost row 0; side 1: from the upper -> row Nx+1)
//   hflag[2 sides]                         sequence number of the last completed halo write per side
//   recs [2 parities][world][8]            one folded record (Rec5, padded) per rank
//   rflag[world]                           sequence number of the last record written by rank w
//   err                                    [0] set when a wait timed out (a peer died or lagged): carried to every rank with the
//                                          records, the run stops (DevState::comm_err);  +8 bytes: the wait bound in ns
//   x                                      fused exchange: hflag2[2], xseq, xdone (see below)
//   seq  [2]                               exchanges / folds this rank has completed (device-resident, so that
//                                          a captured step can be replayed: no sequence number is a kernel argument)
// A sender stores the data into the receiver's buffer, fences at system scope and then publishes the
// sequence number; the receiver's kernel spins on its own flag (bounded by SPIN_NS) and copies the
// rows into its ghost cells.  Slots alternate with the sequence parity: a slot is rewritten two
// exchanges later, which the sender's own stream order places after the receiver has consumed it.
#pragma once
#include <cuda_runtime.h>
#include <vector>
#include "gpf_types.h"

namespace gpf {

struct SlabCommLayout {
    int world, rank, Ny2;
    size_t off_halo, off_hflag, off_recs, off_rflag, off_err, off_seq, off_ptrs, off_x, bytes;
};

struct SlabComm {
    SlabCommLayout lay;
    unsigned char* local = nullptr;
    std::vector<unsigned char*> peer;      // [world]; peer[rank] == local
    unsigned char** d_peer = nullptr;      // the same table on the device (inside `local`)
    cudaIpcMemHandle_t handle;
    bool connected = false;
    static SlabCommLayout layout(int world, int rank, int Ny2);
    int create(int world, int rank, int Ny);
    int connect(const unsigned char* handles);     // [world][sizeof(cudaIpcMemHandle_t)]
    int destroy();
    int set_timeout(double seconds);                // bound of every device-side wait on a peer (default 30 s)
    cudaGraphExec_t step_graph = nullptr;  // one captured slab step (FvSolver::slab_step)
    int graph_lo = -2, graph_hi = -2;
    const void* graph_owner = nullptr;
    double* graph_recs = nullptr;
    double* halo(int r, int parity, int side) const {
        return reinterpret_cast<double*>(peer[r] + lay.off_halo) + ((size_t)parity * 2 + side) * 3 * lay.Ny2;
    }
    unsigned long long* hflag(int r, int side) const { return reinterpret_cast<unsigned long long*>(peer[r] + lay.off_hflag) + side; }
    double* recs(int r, int parity) const { return reinterpret_cast<double*>(peer[r] + lay.off_recs) + (size_t)parity * lay.world * 8; }
    unsigned long long* rflag(int r) const { return reinterpret_cast<unsigned long long*>(peer[r] + lay.off_rflag); }
    int* err() const { return reinterpret_cast<int*>(local + lay.off_err); }
    unsigned long long* seq(int which) const { return reinterpret_cast<unsigned long long*>(local + lay.off_seq) + which; }
    unsigned int* done(int which) const { return reinterpret_cast<unsigned int*>(local + lay.off_seq + 64) + which; }   // block counters
    size_t halo_par_stride() const { return (size_t)2 * 3 * lay.Ny2; }       // doubles between the two parities
    // ---- fused exchange: the pass kernels store edge rows straight into the neighbours' q buffers ----------------
    //   x block of the buffer: hflag2[2] (raised by the lower / upper neighbour), xseq (exchanges published), xdone
    unsigned long long* hflag2(int r, int side) const { return reinterpret_cast<unsigned long long*>(peer[r] + lay.off_x) + side; }
    unsigned long long* xseq() const { return reinterpret_cast<unsigned long long*>(local + lay.off_x) + 2; }
    unsigned int* xdone() const { return reinterpret_cast<unsigned int*>(local + lay.off_x + 24); }
    bool attached = false;
    bool local_only = false;               // peers are communicators of this process (connect_local): nothing to unmap
    double* peer_q[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
    unsigned long long peer_plane[2] = {0, 0};
    int peer_row[2] = {0, 0};
    int nbr[2] = {-1, -1};                 // neighbour ranks (lo, hi)
    void* peer_map[2] = {nullptr, nullptr};   // IPC mappings of the neighbours' workspaces (closed in destroy)
    int attach(const struct FvSolver& S, const unsigned char* blobs, int lo, int hi);   // [world][GPF_PEER_BLOB_BYTES]
    int attach_local(const struct FvSolver& S, const struct FvSolver* lo_s, int lo, const struct FvSolver* hi_s, int hi);
    int connect_local(SlabComm* const* all);
    int detach();                          // back to the exchange kernels: unmaps the neighbours' q buffers
};

}  // namespace gpf
