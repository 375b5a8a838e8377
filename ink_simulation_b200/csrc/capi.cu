// capi.cu — the C ABI of include/ink_capi.h: context, device memory, phase sequencing of System::solve
// (reference src/inkSolver.cpp:5-9, src/waterSolver.cpp:7-96) over the kernels in *_kernels.cuh / cg.cuh.
// There is deliberately no CPU code path: without a CUDA device ink_create fails.
#include "../../include/ink_capi.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "cg.cuh"
#include "common.cuh"
#include "grid_kernels.cuh"
#include "mg.cuh"
#include "mg_vec.cuh"
#include "mg_tri.cuh"
#include "grid_vec.cuh"
#include "nccl_shim.h"
#include "particle_kernels.cuh"

namespace ink {

static thread_local std::string g_create_error;

#define CK(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) {                                                                         \
            char buf_[512];                                                                              \
            snprintf(buf_, sizeof buf_, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            this->err = buf_;                                                                            \
            return 1;                                                                                    \
        }                                                                                                \
    } while (0)

#define NK(call)                                                                                         \
    do {                                                                                                 \
        ncclResult_t r_ = (call);                                                                        \
        if (r_ != ncclSuccess) {                                                                         \
            char buf_[512];                                                                              \
            snprintf(buf_, sizeof buf_, "%s:%d %s -> %s", __FILE__, __LINE__, #call, nccl().GetErrorString(r_)); \
            this->err = buf_;                                                                            \
            return 4;                                                                                    \
        }                                                                                                \
    } while (0)

// launch with the programmatic-stream-serialization attribute (see INK_PDL in common.cuh); INK_NO_PDL=1 disables it
static const bool g_use_pdl = [] { const char* e = getenv("INK_NO_PDL"); return !(e && atoi(e) != 0); }();
// INK_TEST_FAIL_CAPTURE=1 (tests only): behave as if the runtime refused cudaStreamBeginCapture, to exercise the fallback below
static const bool g_fail_capture = [] { const char* e = getenv("INK_TEST_FAIL_CAPTURE"); return e && atoi(e) != 0; }();
template <class... KArgs, class... Args>
static inline void launch_k(void (*kern)(KArgs...), unsigned int grid, unsigned int block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = g_use_pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kern, KArgs(std::forward<Args>(args))...);
}
// one thread-block cluster of `ctas` CTAs (grid == cluster), with the same programmatic-serialization attribute
template <class... KArgs, class... Args>
static inline void launch_cluster(void (*kern)(KArgs...), unsigned int ctas, unsigned int block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ctas); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = ctas; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = g_use_pdl ? 2 : 1;
    cudaLaunchKernelEx(&cfg, kern, KArgs(std::forward<Args>(args))...);
}
#define LAUNCH(kern, grid, block, ...)                                   \
    do {                                                                 \
        launch_k(kern, (grid), (block), 0, stream, __VA_ARGS__);         \
        ++stats.kernel_launches;                                         \
    } while (0)

// frees a temporary device allocation on every way out of a function (the CK macro returns early on error)
struct DevTmp {
    void* p = nullptr;
    ~DevTmp() { if (p) cudaFree(p); }
    template <class T> T* as() const { return static_cast<T*>(p); }
};

static inline unsigned int nblk(long long n, int b = 256) { return (unsigned int)((n + b - 1) / b); }

struct CtxBase {
    ink_params P;
    std::string err;
    ink_stats stats{};
    cudaStream_t stream = nullptr;   // the ctx's own non-blocking stream (created in init; never the legacy NULL stream)
    virtual ~CtxBase() {}
    virtual int init() = 0;
    virtual int set_particles(int drop, const void* pos, const void* vel, int64_t n, int dt) = 0;
    virtual int set_particles_shard(int drop, const void* pos, const void* vel, int64_t n_local, int dt) = 0;
    virtual int num_drops() const = 0;
    virtual int64_t num_particles(int drop) const = 0;
    virtual int get_particles(int drop, void* pos, void* vel, int dt) = 0;
    virtual int set_velocity(const void* u, const void* v, const void* w, int dt) = 0;
    virtual int get_field(int which, void* u, void* v, void* w, int dt) = 0;
    virtual int get_pressure(void* p, int dt) = 0;
    virtual int set_pressure(const void* p, int dt) = 0;
    virtual int get_warm_start(void* q, int dt) = 0;
    virtual int set_warm_start(const void* q, int dt) = 0;
    virtual int get_mark(uint8_t* m) = 0;
    virtual int get_divergence(void* d, int dt) = 0;
    virtual int sample_velocity(const void* pos, int64_t n, void* out, int dt) = 0;
    virtual int step(double ttnr, double* dt_out) = 0;
    virtual int calc_dt(double ttnr, double* dt_out) = 0;
    virtual int advect(double dt) = 0;
    virtual int forces(double dt) = 0;
    virtual int viscosity(double dt) = 0;
    virtual int pressure_rhs(double dt, void* b, int hd) = 0;
    virtual int pressure(double dt) = 0;
    virtual int pressure_given_p(double dt) = 0;
    virtual int poisson_solve(const void* b, void* p, int hd) = 0;
    virtual int vorticity(double dt) = 0;
    virtual int velocity_field_update(double dt) = 0;
    virtual int particles(double dt) = 0;
    virtual int resort() = 0;
    virtual int prefetch_positions(int drop, const void* pos, int64_t n, int hd) = 0;
    virtual int commit_positions(int drop) = 0;
    virtual int fetch_positions_async(int drop, void* pos, int hd) = 0;
    virtual int sync_io() = 0;
    virtual int phase_ms(float out[8]) = 0;
    virtual int check_finite(int64_t out[2]) = 0;
    virtual int seed_emitter(int drop, const float* radius, int nr, const float* si, const float* ci, int ni, const float* sj, const float* cj, int nj) = 0;
    virtual int seed_box(int drop, const int32_t* rnd, int64_t n) = 0;
    virtual int get_phantom(int32_t* have, double v[3]) = 0;
    virtual int set_phantom(int32_t have, const double v[3]) = 0;
};

template <class Real>
struct Ctx final : CtxBase {
    Grid g{};          // g.j0 / g.j1 = the y-slab of rows this rank owns (whole grid on one GPU)
    Consts<Real> C{};
    // multi-GPU (one process per GPU): NCCL communicator over NVLink; every rank keeps FULL-size grid arrays, owns
    // rows [g.j0, g.j1) and a contiguous index shard of every particle drop
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    std::vector<int> slab_lo;   // slab_lo[r] .. slab_lo[r+1] = rows of rank r
    // MAC grid state (struct Cell, src/system.h:41-50) as SoA; vel[cur] is the committed (old == curr) field
    Real* vel[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};
    int cur = 0;
    Real* curl[3] = {nullptr, nullptr, nullptr};
    Real* curl_n = nullptr;   // |curl| per cell (vectorised K7a/K7b pair, grid_vec.cuh)
    Real *pres = nullptr, *q = nullptr, *rhs = nullptr, *cg_r = nullptr, *cg_d[2] = {nullptr, nullptr}, *cg_w = nullptr;
    bool have_q = false;
    // multigrid preconditioner (mg.cuh): lv[0..lt) run as global-memory kernels, the rest inside k_mg_tail
    std::vector<MgLevelPtrs<Real>> lv;
    std::vector<Real*> mg_alloc;
    int mg_lt = 0;
    MgTailArgs mg_tail{};
    size_t mg_tail_smem = 0;
    bool use_mg = false;
    // INK_MG3_FUSED_DOWN=1: levels >= 1 run k_mg3_down_fused (one launch) instead of the residual + restriction kernel pair.  OFF by
    // default: measured on B200 (profiles/r2_e_*) the one-pass kernel is SLOWER per level (16.3 / 14.3 / 12.3 us against 11.2 / 9.2 /
    // 9.1 us for the pair, traced): ~1000 predicated instructions per thread in a grid too small to hide them
    bool mg_fused_down = false;
    bool mv_tile = false;        // INK_CG_MATVEC_TILE=1: k_cg_dir_matvec_tile (shared-memory staging of the direction) instead of k_cg_dir_matvec
    bool mg_nodone = false;
    int mg_down_tile = 0;        // INK_MG3_DOWN_TILE: 1 = k_mg3_down_tile (residual + restriction in one launch) on levels >= 1, 2 = on every level
    int mg_up_tile = 1;          // INK_MG3_UP_TILE: 0 = k_mg3_up everywhere, 1 = k_mg3_up_tile on levels with >= 2 tiles per SM, 2 = wherever legal
    bool mg_coarsest = true;     // INK_MG3_COARSEST=0: the generic shared-memory tail kernel also when the tail is the coarsest level alone
    bool mg_fuse = true;    // INK_MG3_UNFUSED=1 (debug / A-B): separate residual and full 4x4x4 restriction kernels
    size_t mg_tail_cap = 200 * 1024;   // INK_MG_TAIL_KB: shared-memory budget of the single-block tail (experiments)
    int mg_lc = -1;           // >= 1: levels [mg_lc, mg_lt] run in ONE persistent launch (k_mg3_coarse); -1: per-level kernels
    unsigned int* mg_bar = nullptr;
    int mg_cluster = 0;       // > 0: the coarse levels run as one thread-block cluster of this many CTAs (k_mg3_coarse<Real, true>)
    bool mg_tri = false;   // trilinear V(1,1) (mg_tri.cuh) instead of piecewise-constant V(2,2) (mg.cuh)
    int prev_iters = 0;
    // multi-GPU: overlapping-Schwarz extension of the slab the V-cycle runs on (mg_vec.cuh)
    int ovl = 0, je0 = 0, je1 = 0;        // overlap rows; extended slab [je0, je1)
    int dot0 = 0, dot1 = 0;               // rows whose <r,z> the last smoothing kernel may accumulate (outside the bands)
    Real* d_sw = nullptr;                 // sqrt of this rank's partition-of-unity weight per grid row
    Real* band_save[2] = {nullptr, nullptr};
    Real* band_recv[2] = {nullptr, nullptr};
    CgScal* S = nullptr;
    CgScal* hS = nullptr;  // pinned mirror of the scalar head
    // One PCG iteration (CG pair + V-cycle, ~13-20 kernel nodes with their programmatic-dependency edges) captured once as a
    // CUDA graph per direction-buffer parity and replayed: one host call per iteration instead of one per kernel.  Kernels
    // launched after convergence exit on S->done, so replaying a fixed body is legal.  INK_NO_GRAPH=1 = direct launches.
    cudaGraphExec_t it_graph[2] = {nullptr, nullptr};
    const void* it_graph_x = nullptr;     // the solution vector the cached graphs were captured for
    long long it_graph_nodes = 0;
    bool use_graph = true;
    // multi-GPU: per-iteration exchanges by peer stores over NVLink (protocol: cg.cuh) instead of NCCL calls
    bool p2p = false;
    // INK_TRACE_ITER=k: CUDA-event timeline of CG iteration k of every solve, printed by rank 0 (diagnostic)
    int trace_iter = -1; bool tracing = false; bool host_timing = false;
    std::vector<cudaEvent_t> tr_ev; std::vector<const char*> tr_name;
    void TR(const char* name) {
        if (!tracing) return;
        cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, stream); tr_ev.push_back(e); tr_name.push_back(name);
    }
    void tr_dump() {
        if (tr_ev.empty()) return;
        cudaStreamSynchronize(stream);
        if (rank == 0) {
            fprintf(stderr, "[inkb200 trace] iteration %d:", trace_iter);
            for (size_t a = 1; a < tr_ev.size(); ++a) { float ms = 0; cudaEventElapsedTime(&ms, tr_ev[a - 1], tr_ev[a]); fprintf(stderr, " %s %.1f us |", tr_name[a], ms * 1e3f); }
            fprintf(stderr, "\n");
        }
        for (auto e : tr_ev) cudaEventDestroy(e);
        tr_ev.clear(); tr_name.clear();
    }
    unsigned char* arena = nullptr;
    P2pPeers peers{};
    size_t off_halo[2] = {0, 0}, off_band[2] = {0, 0};   // payload regions: [0] written by rank-1, [1] by rank+1
    unsigned long long p2p_seq[2] = {0, 0};
    unsigned long long r_seq = 0;   // sequence number under which the producer of the current residual pushed its boundary rows              // host sequence numbers of channel 0 (r halo) and 1 (z band)
    uint8_t *seed = nullptr, *mark = nullptr;
    unsigned int* d_scal = nullptr;  // [0] max speed bits, [1] oob particle count
    unsigned int* h_scal = nullptr;  // pinned
    bool maxv_valid = false, seed_valid = false, seed_global = false;
    // phantom map cells created by out-of-grid particles (Q8): all carry the same accumulated dt*g
    bool have_phantom = false, oob_now = false;
    Real phantom_v[3] = {0, 0, 0};
    // particles (m_ink, src/system.h:115): all drops concatenated, binned by cell
    Particles<Real> part[2]{};
    int pcur = 0;
    long long pcap = 0, np = 0;
    std::vector<long long> drop_off, drop_n;
    uint32_t *bin_count = nullptr, *bin_key = nullptr, *bin_rank = nullptr;
    std::vector<uint32_t*> scan_lvl;  // tile sums per level
    std::vector<long long> scan_n;
    long long since_sort = 0;
    // K8 tile path (particle_kernels.cuh: TMA-staged shared-memory tiles): work list of slot chunks, rebuilt at every re-binning
    uint2* tile_work = nullptr;
    unsigned int* tile_nwork = nullptr;
    long long tile_work_cap = 0;
    int tile_chunk = 0;        // > 0: the tile kernel runs with chunks of at most this many slots
    // INK_K8_TILE: 0 off (default), 1 automatic (dense clouds, 16-byte rows), 2 forced wherever it is legal.  OFF by default:
    // measured on B200 (profiles/r2_d_*): bit-identical results, but 0.38 ms against 0.24 ms per launch at 10 M particles /
    // 128x256x128 — a warp's 32 particles span ~53 z-cells of one quarter column, so a shared-memory gather still costs 2.1
    // wavefronts (bank conflicts between k and k + 32) against 2.6 L1 wavefronts from global memory: 36.5 M vs 39 M in total, no
    // real saving in the data pipe, while the staging, the bounding-box pass and 3 blocks per SM (56 KB tiles) cost more.
    int tile_mode = 0;
    // staging for host <-> device conversion
    void* stage = nullptr;
    size_t stage_bytes = 0;
    cudaEvent_t ev[10]{};
    bool ev_valid = false;
    // streaming particle I/O (ink_prefetch_positions / ink_commit_positions / ink_fetch_positions_async): two copy streams
    // beside the compute stream, double-buffered device staging, so H2D of the next input and D2H of the last result
    // overlap the step in between
    cudaStream_t s_in = nullptr, s_out = nullptr;
    void* io_in[2] = {nullptr, nullptr};
    void* io_out[2] = {nullptr, nullptr};
    size_t io_in_bytes = 0, io_out_bytes = 0;
    cudaEvent_t ev_in_ready[2]{}, ev_in_free[2]{}, ev_out_ready[2]{}, ev_out_done[2]{};
    int in_next = 0, out_next = 0, in_pending = -1, in_pending_drop = -1, in_pending_hd = 32;

    ~Ctx() override {
        if (stream) cudaStreamSynchronize(stream);
        for (auto& ge : it_graph) if (ge) cudaGraphExecDestroy(ge);
        for (int r = 0; r < world && r < kP2pMaxRanks; ++r)
            if (p2p && r != rank && peers.base[r]) cudaIpcCloseMemHandle(peers.base[r]);
        cudaFree(arena);
        if (comm) nccl().CommDestroy(comm);
        for (auto& b : vel)
            for (auto& p : b) cudaFree(p);
        for (auto& p : curl) cudaFree(p);
        cudaFree(curl_n);
        cudaFree(pres); cudaFree(q); cudaFree(rhs); cudaFree(cg_r); cudaFree(cg_d[0]); cudaFree(cg_d[1]); cudaFree(cg_w);
        for (auto p : mg_alloc) cudaFree(p);
        cudaFree(mg_bar);
        cudaFree(d_sw);
        for (auto p : band_save) cudaFree(p);
        for (auto p : band_recv) cudaFree(p);
        cudaFree(tile_nwork);
        cudaFree(S); cudaFreeHost(hS); cudaFree(seed); cudaFree(mark); cudaFree(d_scal); cudaFreeHost(h_scal);
        free_particles();
        cudaFree(bin_count);
        for (auto p : scan_lvl) cudaFree(p);
        cudaFree(stage);
        for (auto& e : ev)
            if (e) cudaEventDestroy(e);
        for (int b = 0; b < 2; ++b) {
            cudaFree(io_in[b]); cudaFree(io_out[b]);
            for (cudaEvent_t e : {ev_in_ready[b], ev_in_free[b], ev_out_ready[b], ev_out_done[b]})
                if (e) cudaEventDestroy(e);
        }
        if (s_in) cudaStreamDestroy(s_in);
        if (s_out) cudaStreamDestroy(s_out);
        if (stream) cudaStreamDestroy(stream);
    }
    void free_particles() {
        for (auto& pp : part) {
            for (auto& p : pp.p) { cudaFree(p); p = nullptr; }
            for (auto& p : pp.v) { cudaFree(p); p = nullptr; }
            cudaFree(pp.orig); pp.orig = nullptr;
        }
        cudaFree(bin_key); cudaFree(bin_rank);
        bin_key = bin_rank = nullptr;
        cudaFree(tile_work); tile_work = nullptr; tile_work_cap = 0;
        pcap = 0;
    }

    int init() override {
        const ink_params& p = P;
        if (p.X <= 0 || p.Y <= 0 || p.Z <= 0) { err = "grid dims must be positive"; return 2; }
        if ((long long)p.X * p.Y * p.Z >= (1ll << 31)) { err = "grid too large for 32-bit cell indices"; return 2; }
        if (p.world_size < 1 || p.rank < 0 || p.rank >= p.world_size) { err = "bad rank / world_size"; return 2; }
        rank = p.rank; world = p.world_size;
        slab_lo.resize(world + 1);
        for (int r = 0; r <= world; ++r) slab_lo[r] = (int)((long long)p.Y * r / world);
        if (world > 1 && p.Y / world < 4) { err = "y-slab decomposition needs at least 4 grid rows per rank"; return 2; }
        CK(cudaSetDevice(p.device));
        // Own non-blocking stream: the library neither serialises against nor is serialised by the legacy NULL stream (a host
        // process with a blocking stream of its own would otherwise interleave with every one of the step's launches).
        CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        cudaDeviceProp prop;
        CK(cudaGetDeviceProperties(&prop, p.device));
        if (prop.major < 10) { err = "libinkb200 is built for sm_100a (B200) only"; return 3; }
        g.X = p.X; g.Y = p.Y; g.Z = p.Z;
        g.sx = p.Z; g.sy = (long long)p.X * p.Z; g.n = (long long)p.X * p.Y * p.Z;
        g.j0 = slab_lo[rank]; g.j1 = slab_lo[rank + 1];
        if (world > 1) {
            const std::string e = nccl().load();
            if (!e.empty()) { err = e; return 4; }
            ncclUniqueId id;
            static_assert(sizeof(id) <= sizeof(p.nccl_id), "ncclUniqueId does not fit ink_params::nccl_id");
            memcpy(&id, p.nccl_id, sizeof id);
            NK(nccl().CommInitRank(&comm, world, id, rank));
        }
        // constants converted the way the reference's expressions convert them (float first, then Real)
        C.h = (Real)p.cell_dim;
        C.bx = (Real)(float)(p.X * p.cell_dim); C.by = (Real)(float)(p.Y * p.cell_dim); C.bz = (Real)(float)(p.Z * p.cell_dim);
        C.rho = (Real)p.density; C.kvort = (Real)p.k_vort; C.visc = (Real)p.viscosity; C.patm = (Real)p.p_atm;
        for (int a = 0; a < 3; ++a) C.g[a] = (Real)p.gravity[a];
        C.lo = (Real)0.01f;
        C.cx = (Real)(float)(p.X * p.cell_dim - 0.01f); C.cy = (Real)(float)(p.Y * p.cell_dim - 0.01f); C.cz = (Real)(float)(p.Z * p.cell_dim - 0.01f);
        C.unit_h = (p.cell_dim == 1.0f);
        C.R = p.buffer_size;
        const size_t nb = (size_t)g.n * sizeof(Real);
        for (auto& b : vel)
            for (auto& ptr : b) { CK(cudaMalloc(&ptr, nb)); CK(cudaMemsetAsync(ptr, 0, nb, stream)); }
        for (auto& ptr : curl) { CK(cudaMalloc(&ptr, nb)); CK(cudaMemsetAsync(ptr, 0, nb, stream)); }
        CK(cudaMalloc(&curl_n, nb)); CK(cudaMemsetAsync(curl_n, 0, nb, stream));
        for (Real** ptr : {&pres, &q, &rhs, &cg_r, &cg_d[0], &cg_d[1], &cg_w}) { CK(cudaMalloc(ptr, nb)); CK(cudaMemsetAsync(*ptr, 0, nb, stream)); }
        CK(cudaMalloc(&S, sizeof(CgScal))); CK(cudaMemsetAsync(S, 0, sizeof(CgScal), stream));
        { const int dist = world > 1; CK(cudaMemcpyAsync(&S->dist, &dist, sizeof(int), cudaMemcpyHostToDevice, stream)); }
        { const double tol2 = P.cg_rtol * P.cg_rtol; CK(cudaMemcpyAsync(&S->tol2, &tol2, sizeof(double), cudaMemcpyHostToDevice, stream)); CK(cudaStreamSynchronize(stream)); }
        CK(cudaMallocHost(&hS, sizeof(CgScal)));
        CK(cudaMalloc(&seed, g.n)); CK(cudaMemsetAsync(seed, 0, g.n, stream));
        CK(cudaMalloc(&mark, g.n)); CK(cudaMemsetAsync(mark, 0, g.n, stream));
        CK(cudaMalloc(&d_scal, 8 * sizeof(unsigned int))); CK(cudaMemsetAsync(d_scal, 0, 8 * sizeof(unsigned int), stream));   // [4..8): check_finite counters
        CK(cudaMallocHost(&h_scal, 8 * sizeof(unsigned int)));
        // counting-sort scratch: per-cell counts + scan levels
        if ((long long)kBinsPerCell * g.n >= (1ll << 32)) { err = "grid too large for 32-bit particle bin keys"; return 2; }
        CK(cudaMalloc(&bin_count, (size_t)kBinsPerCell * g.n * sizeof(uint32_t)));
        long long m = kBinsPerCell * g.n;
        while (m > 1) {
            m = (m + kScanTile - 1) / kScanTile;
            uint32_t* lvl;
            CK(cudaMalloc(&lvl, (size_t)std::max<long long>(m, 1) * sizeof(uint32_t)));
            scan_lvl.push_back(lvl);
            scan_n.push_back(m);
        }
        for (auto& e : ev) CK(cudaEventCreate(&e));
        if (const char* e = getenv("INK_CG_MATVEC_TILE")) mv_tile = atoi(e) != 0;
        CK(cudaFuncSetAttribute((k_cg_dir_matvec_tile<Real>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cg_matvec_tile_smem<Real>()));
        if (const char* e = getenv("INK_K8_TILE")) tile_mode = atoi(e);
        CK(cudaMalloc(&tile_nwork, sizeof(unsigned int)));
        CK(cudaFuncSetAttribute((k_particles_rk2_tile<Real, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, kTileSmemBytes));
        CK(cudaFuncSetAttribute((k_particles_rk2_tile<Real, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, kTileSmemBytes));
        if (P.cg_precond == INK_PCG_MG || P.cg_precond == INK_PCG_MG_TRI) {
            mg_tri = (P.cg_precond == INK_PCG_MG_TRI);
            if (int rc = mg_setup()) return rc;
        }
        if (world > 1)
            if (int rc = p2p_setup()) return rc;
        maxv_valid = true;  // zero field: max speed bits already 0
        seed_valid = true;
        CK(cudaStreamSynchronize(stream));
        return 0;
    }

    // ---- peer-memory exchanges (cg.cuh) -------------------------------------------------------------
    // One arena per rank (mailboxes, flags, boundary-band payloads), exported over CUDA IPC; the handles travel once through
    // ncclAllGather.  Any failure (no peer access, odd row size, INK_NO_P2P=1) leaves p2p == false: the NCCL path is kept.
    int p2p_setup() {
        if (const char* e = getenv("INK_NO_P2P")) if (atoi(e) != 0) return 0;
        if (const char* e = getenv("INK_NO_FOLD")) fold_ar = (atoi(e) == 0);
        if (world > kP2pMaxRanks) return 0;
        const size_t rowb = (size_t)g.sy * sizeof(Real);
        if (rowb % 16 != 0) return 0;
        const size_t head = (sizeof(P2pHead) + 255) / 256 * 256;
        const size_t halo = (size_t)ovl * rowb, band = (size_t)2 * ovl * rowb;
        off_halo[0] = head; off_halo[1] = head + halo;
        off_band[0] = head + 2 * halo; off_band[1] = head + 2 * halo + band;
        const size_t bytes = head + 2 * halo + 2 * band;
        CK(cudaMalloc(&arena, bytes));
        CK(cudaMemsetAsync(arena, 0, bytes, stream));
        CK(cudaStreamSynchronize(stream));
        cudaIpcMemHandle_t mine;
        CK(cudaIpcGetMemHandle(&mine, arena));
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
        unsigned char* hbuf = nullptr;
        CK(cudaMalloc(&hbuf, (size_t)world * 64));
        CK(cudaMemcpyAsync(hbuf + (size_t)rank * 64, &mine, 64, cudaMemcpyHostToDevice, stream));
        NK(nccl().AllGather(hbuf + (size_t)rank * 64, hbuf, 64, ncclUint8, comm, stream));
        CK(cudaStreamSynchronize(stream));
        std::vector<cudaIpcMemHandle_t> all((size_t)world);
        CK(cudaMemcpyAsync(all.data(), hbuf, (size_t)world * 64, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        cudaFree(hbuf);
        int ok = 1;
        for (int r = 0; r < world; ++r) {
            if (r == rank) { peers.base[r] = arena; continue; }
            void* ptr = nullptr;
            if (cudaIpcOpenMemHandle(&ptr, all[(size_t)r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = 0; break; }
            peers.base[r] = (unsigned char*)ptr;
        }
        // all ranks must agree (a rank that could not map a peer forces everybody back to NCCL)
        int* dok = nullptr;
        CK(cudaMalloc(&dok, sizeof(int)));
        CK(cudaMemcpyAsync(dok, &ok, sizeof(int), cudaMemcpyHostToDevice, stream));
        NK(nccl().AllReduce(dok, dok, 1, ncclInt32, ncclMin, comm, stream));
        CK(cudaStreamSynchronize(stream));
        CK(cudaMemcpyAsync(&ok, dok, sizeof(int), cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        cudaFree(dok);
        p2p = ok != 0;
        if (rank == 0 && getenv("INK_VERBOSE")) fprintf(stderr, "[inkb200] world %d: per-iteration exchanges over %s\n", world, p2p ? "peer memory (CUDA IPC + NVLink stores)" : "NCCL");
        return 0;
    }
    P2pHead* head_of(int r) const { return reinterpret_cast<P2pHead*>(peers.base[r]); }
    // sum S->loc[off .. off+count) over the ranks, then the scalar epilogue `stage` of k_cg_post
    int ar_post(int off, int count, int stage) {
        if (p2p) {
            launch_k(k_p2p_allreduce_post, 1, 32, 0, stream, S, peers, rank, world, off, count, stage, use_mg ? 1 : 0);
            ++stats.kernel_launches;
            return 0;
        }
        if (int rc = allreduce(S->loc + off, (size_t)count, ncclDouble, ncclSum)) return rc;
        return cg_post(stage);
    }
    // the scalar all-reduce as the epilogue of the reducing kernel's last block (peer-memory path only; INK_NO_FOLD=1 = separate launch)
    bool fold_ar = true;
    ArFold make_fold(int off, int count, int stage) {
        ArFold f{};
        if (!(p2p && fold_ar)) return f;
        f.on = 1; f.rank = rank; f.world = world; f.off = off; f.count = count; f.stage = stage; f.precond = use_mg ? 1 : 0;
        f.peers = peers;
        return f;
    }
    // boundary rows of the residual about to be produced (k_cg_init / k_cg_update) go straight into the neighbours' arenas
    P2pPush make_push() {
        P2pPush p{};
        if (!(p2p && use_mg)) return p;
        p.on = 1;
        p.seq = r_seq = ++p2p_seq[0];
        p.lo_cells = (long long)ovl * g.sy;
        p.hi_start = (long long)(g.j1 - g.j0 - ovl) * g.sy;
        if (rank > 0) { p.dst_lo = peers.base[rank - 1] + off_halo[1]; p.flag_lo = &head_of(rank - 1)->flag[0][1]; }
        if (rank + 1 < world) { p.dst_hi = peers.base[rank + 1] + off_halo[0]; p.flag_hi = &head_of(rank + 1)->flag[0][0]; }
        return p;
    }

    // ---- helpers ---------------------------------------------------------------------------------
    // device -> pinned host, a few words, by a kernel instead of the D2H copy engine (see k_copy_words)
    void read_small(void* host_pinned, const void* dev, size_t bytes) {
        launch_k(k_copy_words, 1, 64, 0, stream, (uint32_t*)host_pinned, (const uint32_t*)dev, (int)(bytes / 4));
        ++stats.kernel_launches;
    }
    static bool is_device_ptr(const void* p) {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
        return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
    }
    int need_stage(size_t bytes) {
        if (bytes <= stage_bytes) return 0;
        if (stage) CK(cudaFree(stage));
        stage = nullptr; stage_bytes = 0;
        CK(cudaMalloc(&stage, bytes));
        stage_bytes = bytes;
        return 0;
    }
    // bring `bytes` of caller memory onto the device (no copy if it already is device memory)
    int to_device(const void* src, size_t bytes, size_t stage_off, const void** out) {
        if (is_device_ptr(src)) { *out = src; return 0; }
        CK(cudaMemcpyAsync((char*)stage + stage_off, src, bytes, cudaMemcpyHostToDevice, stream));
        *out = (char*)stage + stage_off;
        return 0;
    }
    Real* U(int a) { return vel[cur][a]; }
    Real* Un(int a) { return vel[cur ^ 1][a]; }
    unsigned int cell_blocks() const { return nblk((long long)(g.j1 - g.j0) * g.sy); }
    static unsigned int cell_blocks(const Grid& q) { return nblk((long long)(q.j1 - q.j0) * q.sy); }
    // vectorised stencil kernels (grid_vec.cuh): one thread per 4 z-consecutive cells
    static unsigned int vec_blocks(const Grid& q) { return nblk((long long)(q.j1 - q.j0) * q.X * (q.Z / 4)); }
    // this rank's slab widened by `ext` rows each side (clipped to the grid): phases before a stencil phase are computed
    // redundantly on the rows the stencil will read, so the velocity phases need NO halo exchange (multi-GPU)
    Grid slab(int ext) const {
        Grid q = g;
        if (world > 1) { q.j0 = std::max(0, g.j0 - ext); q.j1 = std::min(g.Y, g.j1 + ext); }
        return q;
    }
    static ncclDataType_t nc_real() { return sizeof(Real) == 4 ? ncclFloat : ncclDouble; }
    // rows [j0-nlo, j0) <- rank-1's top rows, rows [j1, j1+nhi) <- rank+1's bottom rows (array indexed like the grid)
    int halo_exchange(Real* a, int nlo, int nhi) {
        if (world == 1) return 0;
        const size_t row = (size_t)g.sy;
        NK(nccl().GroupStart());
        if (rank > 0) {
            if (nlo) NK(nccl().Recv(a + (size_t)(g.j0 - nlo) * row, nlo * row, nc_real(), rank - 1, comm, stream));
            if (nhi) NK(nccl().Send(a + (size_t)g.j0 * row, nhi * row, nc_real(), rank - 1, comm, stream));
        }
        if (rank + 1 < world) {
            if (nhi) NK(nccl().Recv(a + (size_t)g.j1 * row, nhi * row, nc_real(), rank + 1, comm, stream));
            if (nlo) NK(nccl().Send(a + (size_t)(g.j1 - nlo) * row, nlo * row, nc_real(), rank + 1, comm, stream));
        }
        NK(nccl().GroupEnd());
        return 0;
    }
    // every rank's own rows -> every rank (in place): one broadcast per owner inside a group
    int allgather_rows(void* a, size_t elem) {
        if (world == 1) return 0;
        const size_t row = (size_t)g.sy * elem;
        if (g.Y % world == 0) {   // equal slabs: one in-place all-gather
            const size_t cnt = (size_t)(g.Y / world) * row;
            NK(nccl().AllGather((char*)a + (size_t)rank * cnt, a, cnt, ncclChar, comm, stream));
            return 0;
        }
        NK(nccl().GroupStart());
        for (int r = 0; r < world; ++r) {
            char* ptr = (char*)a + (size_t)slab_lo[r] * row;
            NK(nccl().Broadcast(ptr, ptr, (size_t)(slab_lo[r + 1] - slab_lo[r]) * row, ncclChar, r, comm, stream));
        }
        NK(nccl().GroupEnd());
        return 0;
    }
    int allgather_velocity() {   // the three components as ONE grouped NCCL operation (one launch, one pass of the protocol)
        if (world == 1) return 0;
        NK(nccl().GroupStart());
        for (int a = 0; a < 3; ++a)
            if (int rc = allgather_rows(U(a), sizeof(Real))) { nccl().GroupEnd(); return rc; }
        NK(nccl().GroupEnd());
        return 0;
    }
    int allreduce(void* buf, size_t count, ncclDataType_t t, ncclRedOp_t op) {
        if (world == 1) return 0;
        NK(nccl().AllReduce(buf, buf, count, t, op, comm, stream));
        return 0;
    }
    bool vec4() const { return (g.Z % 4) == 0; }
    unsigned int cg_blocks() const {
        const long long nvec = (long long)(g.j1 - g.j0) * g.X * (g.Z / (vec4() ? 4 : 1));
        return (unsigned int)std::min<long long>(kCgBlocks, std::max<long long>(1, (nvec + 255) / 256));
    }

    template <class T>
    int grid_in(const void* src, Real* dst) {   // public (X,Y,Z) array of T -> internal
        const void* d;
        if (int rc = to_device(src, (size_t)g.n * sizeof(T), 0, &d)) return rc;
        LAUNCH((k_xyz_to_internal<T, Real>), nblk(g.n), 256, g, (const T*)d, dst);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(stream));  // staging is reused by the next call; a caller's device buffer is free again on return
        return 0;
    }
    template <class T>
    int grid_out(const Real* src, void* dst) {
        if (is_device_ptr(dst)) {
            LAUNCH((k_internal_to_xyz<Real, T>), nblk(g.n), 256, g, src, (T*)dst);
            CK(cudaGetLastError());
            CK(cudaStreamSynchronize(stream));   // device outputs are complete when the call returns (the ctx stream is private)
            return 0;
        }
        LAUNCH((k_internal_to_xyz<Real, T>), nblk(g.n), 256, g, src, (T*)stage);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(dst, stage, (size_t)g.n * sizeof(T), cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        return 0;
    }
    int grid_in_any(const void* src, Real* dst, int hd) {
        if (int rc = need_stage((size_t)g.n * 8)) return rc;
        return hd == 64 ? grid_in<double>(src, dst) : grid_in<float>(src, dst);
    }
    int grid_out_any(const Real* src, void* dst, int hd) {
        if (int rc = need_stage((size_t)g.n * 8)) return rc;
        return hd == 64 ? grid_out<double>(src, dst) : grid_out<float>(src, dst);
    }
    static int check_hd(int hd, std::string& err) {
        if (hd != 32 && hd != 64) { err = "host_dtype must be 32 or 64"; return 2; }
        return 0;
    }

    // ---- particles ---------------------------------------------------------------------------------
    int reserve_particles(long long n) {
        if (n <= pcap) return 0;
        // keep existing content
        Particles<Real> old[2] = {part[0], part[1]};
        uint32_t *okey = bin_key, *orank = bin_rank;
        const long long ncap = std::max<long long>(n, 1);
        for (int b = 0; b < 2; ++b) {
            for (int a = 0; a < 3; ++a) { CK(cudaMalloc(&part[b].p[a], (size_t)ncap * sizeof(Real))); CK(cudaMalloc(&part[b].v[a], (size_t)ncap * sizeof(Real))); }
            CK(cudaMalloc(&part[b].orig, (size_t)ncap * sizeof(uint32_t)));
        }
        CK(cudaMalloc(&bin_key, (size_t)ncap * sizeof(uint32_t)));
        CK(cudaMalloc(&bin_rank, (size_t)ncap * sizeof(uint32_t)));
        if (np > 0) {
            for (int a = 0; a < 3; ++a) {
                CK(cudaMemcpyAsync(part[pcur].p[a], old[pcur].p[a], (size_t)np * sizeof(Real), cudaMemcpyDeviceToDevice, stream));
                CK(cudaMemcpyAsync(part[pcur].v[a], old[pcur].v[a], (size_t)np * sizeof(Real), cudaMemcpyDeviceToDevice, stream));
            }
            CK(cudaMemcpyAsync(part[pcur].orig, old[pcur].orig, (size_t)np * sizeof(uint32_t), cudaMemcpyDeviceToDevice, stream));
        }
        CK(cudaStreamSynchronize(stream));
        for (auto& o : old) {
            for (auto& p : o.p) cudaFree(p);
            for (auto& p : o.v) cudaFree(p);
            cudaFree(o.orig);
        }
        cudaFree(okey); cudaFree(orank);
        pcap = ncap;
        return 0;
    }

    // particles are stored in caller order first (orig = identity over the concatenated drops), then binned
    int set_particles(int drop, const void* pos, const void* vel_, int64_t n, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        if (drop < 0 || n < 0) { err = "bad drop / n"; return 2; }
        if (world > 1) {   // every rank is handed the full drop and keeps the contiguous index shard [n*r/W, n*(r+1)/W)
            const long long first = (long long)n * rank / world, last = (long long)n * (rank + 1) / world;
            const size_t es0 = hd == 64 ? 8 : 4;
            pos = (const char*)pos + (size_t)first * 3 * es0;
            if (vel_) vel_ = (const char*)vel_ + (size_t)first * 3 * es0;
            n = last - first;
        }
        return set_particles_shard(drop, pos, vel_, n, hd);
    }
    // this rank's particles of the drop only (what a per-rank loader would hand over); single GPU: the whole drop
    int set_particles_shard(int drop, const void* pos, const void* vel_, int64_t n, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        if (drop < 0 || n < 0) { err = "bad drop / n"; return 2; }
        if (drop > (int)drop_n.size()) { err = "drops must be added in order (0, 1, ...)"; return 2; }
        // rebuild the concatenation in caller order: un-bin what we have, replace / append this drop
        std::vector<long long> new_n = drop_n;
        if (drop == (int)new_n.size()) new_n.push_back(n); else new_n[drop] = n;
        std::vector<long long> new_off(new_n.size());
        long long tot = 0;
        for (size_t d = 0; d < new_n.size(); ++d) { new_off[d] = tot; tot += new_n[d]; }
        if (tot >= (1ll << 32)) { err = "more than 2^32-1 particles on one device"; return 2; }
        const size_t es = hd == 64 ? 8 : 4;
        // 1) export the other drops (caller order) into a temporary AoS snapshot in Real
        std::vector<Real*> keep_pos(drop_n.size(), nullptr), keep_vel(drop_n.size(), nullptr);
        for (size_t d = 0; d < drop_n.size(); ++d) {
            if ((int)d == drop || drop_n[d] == 0) continue;
            CK(cudaMalloc(&keep_pos[d], (size_t)drop_n[d] * 3 * sizeof(Real)));
            CK(cudaMalloc(&keep_vel[d], (size_t)drop_n[d] * 3 * sizeof(Real)));
            LAUNCH((k_export_aos<Real, Real>), nblk(np), 256, np, drop_off[d], drop_n[d], part[pcur], keep_pos[d], keep_vel[d]);
        }
        CK(cudaGetLastError());
        const long long old_np = np;
        np = 0;  // content is rebuilt below; do not copy in reserve
        if (int rc = reserve_particles(tot)) { np = old_np; return rc; }
        for (size_t d = 0; d < new_n.size(); ++d) {
            if (new_n[d] == 0) continue;
            if ((int)d == drop) {
                if (int rc = need_stage((size_t)n * 3 * es * 2)) return rc;
                const void *dp, *dv = nullptr;
                if (int rc = to_device(pos, (size_t)n * 3 * es, 0, &dp)) return rc;
                if (vel_) if (int rc = to_device(vel_, (size_t)n * 3 * es, (size_t)n * 3 * es, &dv)) return rc;
                if (hd == 64) LAUNCH((k_import_aos<double, Real>), nblk(n), 256, n, new_off[d], (const double*)dp, (const double*)dv, part[pcur]);
                else LAUNCH((k_import_aos<float, Real>), nblk(n), 256, n, new_off[d], (const float*)dp, (const float*)dv, part[pcur]);
            } else {
                LAUNCH((k_import_aos<Real, Real>), nblk(new_n[d]), 256, new_n[d], new_off[d], (const Real*)keep_pos[d], (const Real*)keep_vel[d], part[pcur]);
            }
        }
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(stream));
        for (auto p : keep_pos) cudaFree(p);
        for (auto p : keep_vel) cudaFree(p);
        drop_n = new_n; drop_off = new_off; np = tot;
        seed_valid = false;
        return resort();
    }
    // ---- seeding on the device: the arrays are produced by a kernel, then handed to set_particles_shard as
    // device pointers (no PCIe traffic beyond the tables)
    int seed_emitter(int drop, const float* radius, int nr, const float* si, const float* ci, int ni, const float* sj, const float* cj, int nj) override {
        if (nr <= 0 || ni <= 0 || nj <= 0) { err = "seed_emitter: empty tables"; return 2; }
        const long long n = (long long)nr * ni * nj;
        const long long first = n * rank / world, count = n * (rank + 1) / world - first;
        const size_t ntab = (size_t)nr + 2 * (size_t)ni + 2 * (size_t)nj;
        DevTmp t_tab, t_pos, t_vel;   // freed on every return, the early error returns of CK included
        CK(cudaMalloc(&t_tab.p, ntab * sizeof(float)));
        CK(cudaMalloc(&t_pos.p, (size_t)std::max<long long>(count, 1) * 3 * sizeof(Real)));
        CK(cudaMalloc(&t_vel.p, (size_t)std::max<long long>(count, 1) * 3 * sizeof(Real)));
        float* tab = t_tab.as<float>();
        Real *dpos = t_pos.as<Real>(), *dvel = t_vel.as<Real>();
        float *tr = tab, *tsi = tr + nr, *tci = tsi + ni, *tsj = tci + ni, *tcj = tsj + nj;
        CK(cudaMemcpyAsync(tr, radius, nr * sizeof(float), cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(tsi, si, ni * sizeof(float), cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(tci, ci, ni * sizeof(float), cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(tsj, sj, nj * sizeof(float), cudaMemcpyHostToDevice, stream));
        CK(cudaMemcpyAsync(tcj, cj, nj * sizeof(float), cudaMemcpyHostToDevice, stream));
        const float c = std::cos((float)M_PI);   // the only surviving term of the degenerate rotation (src/inkSolver.cpp:16-18)
        if (count > 0)
            LAUNCH((k_seed_emitter<Real>), nblk(count), 256, first, count, ni, nj, (const float*)tr, (const float*)tsi, (const float*)tci, (const float*)tsj,
                   (const float*)tcj, c, P.X / 2.f, P.Y - 0.1f, P.Z / 2.f, dpos, dvel);
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(stream));
        return set_particles_shard(drop, dpos, dvel, count, (int)sizeof(Real) * 8);
    }
    int seed_box(int drop, const int32_t* rnd, int64_t n) override {
        if (n < 0) { err = "seed_box: bad n"; return 2; }
        const long long first = (long long)n * rank / world, count = (long long)n * (rank + 1) / world - first;
        DevTmp t_rnd, t_pos, t_vel;
        CK(cudaMalloc(&t_rnd.p, (size_t)std::max<long long>(count, 1) * 3 * sizeof(int32_t)));
        CK(cudaMalloc(&t_pos.p, (size_t)std::max<long long>(count, 1) * 3 * sizeof(Real)));
        CK(cudaMalloc(&t_vel.p, (size_t)std::max<long long>(count, 1) * 3 * sizeof(Real)));
        int32_t* drnd = t_rnd.as<int32_t>();
        Real *dpos = t_pos.as<Real>(), *dvel = t_vel.as<Real>();
        if (count > 0) {
            CK(cudaMemcpyAsync(drnd, rnd + 3 * first, (size_t)count * 3 * sizeof(int32_t), cudaMemcpyHostToDevice, stream));
            // bounds exactly as the reference writes them: float expressions for x and z, double literals for y (src/system.cpp:102-104)
            LAUNCH((k_seed_box<Real>), nblk(count), 256, count, (const int32_t*)drnd, (double)(P.X / 2.f - P.X / 5.f), (double)(P.X / 2.f + P.X / 5.f),
                   P.Y - 1.5, P.Y - 1., (double)(P.Z / 2.f - P.Z / 3.f), (double)(P.Z / 2.f + P.Z / 3.f), dpos, dvel);
        }
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(stream));
        return set_particles_shard(drop, dpos, dvel, count, (int)sizeof(Real) * 8);
    }
    int num_drops() const override { return (int)drop_n.size(); }
    int64_t num_particles(int drop) const override { return (drop >= 0 && drop < (int)drop_n.size()) ? drop_n[drop] : 0; }

    int get_particles(int drop, void* pos, void* vel_, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        if (drop < 0 || drop >= (int)drop_n.size()) { err = "no such drop"; return 2; }
        const long long n = drop_n[drop];
        if (n == 0) return 0;
        const size_t es = hd == 64 ? 8 : 4, bytes = (size_t)n * 3 * es;
        const bool dpos = pos && is_device_ptr(pos), dvel = vel_ && is_device_ptr(vel_);
        if (int rc = need_stage(bytes * 2)) return rc;
        void* op = pos ? (dpos ? pos : stage) : nullptr;
        void* ov = vel_ ? (dvel ? vel_ : (char*)stage + bytes) : nullptr;
        if (hd == 64) LAUNCH((k_export_aos<Real, double>), nblk(np), 256, np, drop_off[drop], n, part[pcur], (double*)op, (double*)ov);
        else LAUNCH((k_export_aos<Real, float>), nblk(np), 256, np, drop_off[drop], n, part[pcur], (float*)op, (float*)ov);
        CK(cudaGetLastError());
        if (pos && !dpos) CK(cudaMemcpyAsync(pos, op, bytes, cudaMemcpyDeviceToHost, stream));
        if (vel_ && !dvel) CK(cudaMemcpyAsync(vel_, ov, bytes, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        return 0;
    }

    // ---- streaming particle I/O ----------------------------------------------------------------------------------
    int io_setup() {
        if (s_in) return 0;
        CK(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
        for (int b = 0; b < 2; ++b) {
            CK(cudaEventCreateWithFlags(&ev_in_ready[b], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&ev_in_free[b], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&ev_out_ready[b], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&ev_out_done[b], cudaEventDisableTiming));
        }
        return 0;
    }
    // start the H2D copy of the next positions of `drop` (same particle count as now; pinned host memory makes it truly
    // asynchronous) on the input copy stream; returns at once
    int prefetch_positions(int drop, const void* pos, int64_t n, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        if (drop < 0 || drop >= (int)drop_n.size() || n != drop_n[drop]) { err = "prefetch_positions: drop must exist with the same particle count (use ink_set_particles first)"; return 2; }
        if (in_pending >= 0) { err = "prefetch_positions: a prefetched buffer is waiting for ink_commit_positions"; return 2; }
        if (int rc = io_setup()) return rc;
        const size_t bytes = (size_t)n * 3 * (hd == 64 ? 8 : 4);
        if (bytes > io_in_bytes) {
            CK(cudaDeviceSynchronize());
            for (int b = 0; b < 2; ++b) { cudaFree(io_in[b]); io_in[b] = nullptr; CK(cudaMalloc(&io_in[b], bytes)); }
            io_in_bytes = bytes;
        }
        const int b = in_next;
        CK(cudaStreamWaitEvent(s_in, ev_in_free[b], 0));     // the import kernel that last read this buffer
        if (is_device_ptr(pos)) CK(cudaMemcpyAsync(io_in[b], pos, bytes, cudaMemcpyDeviceToDevice, s_in));
        else CK(cudaMemcpyAsync(io_in[b], pos, bytes, cudaMemcpyHostToDevice, s_in));
        CK(cudaEventRecord(ev_in_ready[b], s_in));
        in_pending = b; in_pending_drop = drop; in_pending_hd = hd;
        in_next ^= 1;
        return 0;
    }
    // make the prefetched positions the state of the drop: the compute stream waits for the copy, then gathers them into the
    // binned slots (binning and velocities kept; the periodic re-binning of ink_step keeps the layout tight)
    int commit_positions(int drop) override {
        if (in_pending < 0 || in_pending_drop != drop) { err = "commit_positions: nothing prefetched for this drop"; return 2; }
        const int b = in_pending;
        CK(cudaStreamWaitEvent(stream, ev_in_ready[b], 0));
        if (in_pending_hd == 64) LAUNCH((k_import_pos_perm<double, Real>), nblk(np), 256, np, drop_off[drop], drop_n[drop], part[pcur], (const double*)io_in[b]);
        else LAUNCH((k_import_pos_perm<float, Real>), nblk(np), 256, np, drop_off[drop], drop_n[drop], part[pcur], (const float*)io_in[b]);
        CK(cudaGetLastError());
        CK(cudaEventRecord(ev_in_free[b], stream));
        in_pending = -1;
        seed_valid = false;
        return 0;
    }
    // positions of `drop` (caller order) -> pos, asynchronously: export on the compute stream, D2H on the output copy
    // stream.  `pos` is complete after ink_sync_io (or after the second next ink_fetch_positions_async returned).
    int fetch_positions_async(int drop, void* pos, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        if (drop < 0 || drop >= (int)drop_n.size()) { err = "no such drop"; return 2; }
        const long long n = drop_n[drop];
        if (n == 0) return 0;
        if (int rc = io_setup()) return rc;
        const size_t bytes = (size_t)n * 3 * (hd == 64 ? 8 : 4);
        if (bytes > io_out_bytes) {
            CK(cudaDeviceSynchronize());
            for (int b = 0; b < 2; ++b) { cudaFree(io_out[b]); io_out[b] = nullptr; CK(cudaMalloc(&io_out[b], bytes)); }
            io_out_bytes = bytes;
        }
        const int b = out_next;
        CK(cudaStreamWaitEvent(stream, ev_out_done[b], 0));   // the D2H that last read this staging buffer
        if (hd == 64) LAUNCH((k_export_aos<Real, double>), nblk(np), 256, np, drop_off[drop], n, part[pcur], (double*)io_out[b], (double*)nullptr);
        else LAUNCH((k_export_aos<Real, float>), nblk(np), 256, np, drop_off[drop], n, part[pcur], (float*)io_out[b], (float*)nullptr);
        CK(cudaGetLastError());
        CK(cudaEventRecord(ev_out_ready[b], stream));
        CK(cudaStreamWaitEvent(s_out, ev_out_ready[b], 0));
        if (is_device_ptr(pos)) CK(cudaMemcpyAsync(pos, io_out[b], bytes, cudaMemcpyDeviceToDevice, s_out));
        else CK(cudaMemcpyAsync(pos, io_out[b], bytes, cudaMemcpyDeviceToHost, s_out));
        CK(cudaEventRecord(ev_out_done[b], s_out));
        out_next ^= 1;
        return 0;
    }
    int sync_io() override {
        if (s_in) CK(cudaStreamSynchronize(s_in));
        CK(cudaStreamSynchronize(stream));
        if (s_out) CK(cudaStreamSynchronize(s_out));
        return 0;
    }

    int scan_counts() {   // exclusive scan of bin_count in place
        long long m = kBinsPerCell * g.n;
        uint32_t* in = bin_count;
        size_t lvl = 0;
        std::vector<uint32_t*> ins;
        std::vector<long long> ns;
        while (true) {
            const long long tiles = (m + kScanTile - 1) / kScanTile;
            LAUNCH(k_scan_tiles, (unsigned int)tiles, 256, m, in, in, scan_lvl[lvl]);
            ins.push_back(in); ns.push_back(m);
            if (tiles == 1) break;
            in = scan_lvl[lvl]; m = tiles; ++lvl;
        }
        for (int l = (int)ins.size() - 2; l >= 0; --l) LAUNCH(k_scan_add, nblk(ns[l]), 256, ns[l], ins[l], ins[l + 1]);
        CK(cudaGetLastError());
        return 0;
    }
    int resort() override {
        since_sort = 0;
        if (np == 0) return 0;
        CK(cudaMemsetAsync(bin_count, 0, (size_t)kBinsPerCell * g.n * sizeof(uint32_t), stream));
        LAUNCH((k_bin_count<Real>), nblk(np), 256, g, np, part[pcur], bin_count, bin_key, bin_rank);
        if (int rc = scan_counts()) return rc;
        LAUNCH((k_bin_scatter<Real>), nblk(np), 256, np, part[pcur], part[pcur ^ 1], bin_count, bin_key, bin_rank);
        CK(cudaGetLastError());
        pcur ^= 1;
        ++stats.resorts;
        return tile_plan();
    }
    // Decide whether K8 runs in its tile form for the current cloud and, if so, cut every half-row's slot range into chunks.
    // Chunk size: the largest multiple of 256 slots (<= 1024) whose expected tile — a chunk of a uniform cloud spans
    // w = chunk / (particles per half cell column) cells in x: 9 w + 21 z-rows — stays under 40 KB (the kernel has 56 KB).
    int tile_plan() {
        tile_chunk = 0;
        if (tile_mode == 0 || np == 0 || !C.unit_h || g.n >= (1ll << 31) || ((size_t)g.Z * sizeof(Real)) % 16 != 0) return 0;
        const double per_half_col = (double)np / (2.0 * g.X * g.Y);
        int chunk = 0;
        for (int ch = 256 * kTileKP; ch >= 256; ch -= 256) {
            const double w = ch / std::max(per_half_col, 1e-9);
            if ((9.0 * w + 21.0) * g.Z * sizeof(Real) <= 40.0 * 1024) { chunk = ch; break; }
        }
        const long long max_work = (np + 255) / 256 + 2ll * g.Y + 1;
        if (tile_mode == 1) {
            if (chunk == 0) return 0;                                         // sparse cloud or very long rows: tiles would not fit
            if ((double)np / ((double)np / chunk + 2.0 * g.Y) < 128.0) return 0;   // too few particles per staged tile to pay for it
        } else if (chunk == 0) chunk = 256;
        if (max_work > tile_work_cap) {
            if (tile_work) CK(cudaFree(tile_work));
            tile_work = nullptr; tile_work_cap = 0;
            CK(cudaMalloc(&tile_work, (size_t)max_work * sizeof(uint2)));
            tile_work_cap = max_work;
        }
        CK(cudaMemsetAsync(tile_nwork, 0, sizeof(unsigned int), stream));
        LAUNCH(k_tile_worklist, (unsigned int)(2 * g.Y), 128, g.X, g.Y, g.Z, np, (const uint32_t*)bin_count, chunk, tile_work, tile_nwork);
        CK(cudaGetLastError());
        tile_chunk = chunk;
        return 0;
    }
    int ensure_seed() {
        if (seed_valid) {
            if (world > 1 && !seed_global) {
                if (int rc = allreduce(seed, (size_t)g.n, ncclUint8, ncclMax)) return rc;
                seed_global = true;
            }
            return 0;
        }
        CK(cudaMemsetAsync(seed, 0, g.n, stream));
        CK(cudaMemsetAsync(d_scal + 1, 0, sizeof(unsigned int), stream));
        if (np > 0) LAUNCH((k_mark_particle_cells<Real>), nblk(np), 256, g, np, part[pcur], seed, d_scal + 1);
        CK(cudaGetLastError());
        if (int rc = allreduce(d_scal + 1, 1, ncclUint32, ncclSum)) return rc;
        if (int rc = allreduce(seed, (size_t)g.n, ncclUint8, ncclMax)) return rc;   // OR of every rank's particle cells
        seed_global = true;
        read_small(h_scal + 1, d_scal + 1, sizeof(unsigned int));
        CK(cudaStreamSynchronize(stream));
        oob_now = h_scal[1] > 0;
        if (oob_now) have_phantom = true;
        seed_valid = true;
        return 0;
    }

    // ---- grid state --------------------------------------------------------------------------------
    int set_velocity(const void* u, const void* v, const void* w, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        const void* src[3] = {u, v, w};
        for (int a = 0; a < 3; ++a)
            if (int rc = grid_in_any(src[a], U(a), hd)) return rc;
        maxv_valid = false;
        return 0;
    }
    int get_field(int which, void* u, void* v, void* w, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        void* dst[3] = {u, v, w};
        for (int a = 0; a < 3; ++a) {
            if (!dst[a]) continue;
            const Real* src = (which == INK_FIELD_CURL) ? curl[a] : U(a);   // old == curr between phases
            if (int rc = grid_out_any(src, dst[a], hd)) return rc;
        }
        return 0;
    }
    int get_pressure(void* p, int hd) override {   // multi-GPU: collective (every rank calls it)
        if (int rc = check_hd(hd, err)) return rc;
        if (int rc = allgather_rows(pres, sizeof(Real))) return rc;
        return grid_out_any(pres, p, hd);
    }
    int set_pressure(const void* p, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        if (int rc = grid_in_any(p, pres, hd)) return rc;
        LAUNCH((k_add_scalar<Real>), nblk(g.n), 256, g.n, -C.patm, pres, q);
        CK(cudaGetLastError());
        have_q = true;
        return 0;
    }
    // the Poisson solver's own unknown q = p - p_atm, exactly as held (checkpoints: p has lost the low bits of q to the +p_atm)
    int get_warm_start(void* qo, int hd) override {   // multi-GPU: collective
        if (int rc = check_hd(hd, err)) return rc;
        if (!have_q) CK(cudaMemsetAsync(q, 0, (size_t)g.n * sizeof(Real), stream));
        if (int rc = allgather_rows(q, sizeof(Real))) return rc;
        return grid_out_any(q, qo, hd);
    }
    int set_warm_start(const void* qi, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        if (int rc = grid_in_any(qi, q, hd)) return rc;
        have_q = true;
        return 0;
    }
    int get_mark(uint8_t* m) override {
        if (int rc = need_stage((size_t)g.n * 8)) return rc;
        uint8_t* dst = is_device_ptr(m) ? m : (uint8_t*)stage;
        if (int rc = allgather_rows(mark, 1)) return rc;
        LAUNCH(k_mark_to_xyz, nblk(g.n), 256, g, mark, dst);
        CK(cudaGetLastError());
        if (dst != m) CK(cudaMemcpyAsync(m, dst, g.n, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        return 0;
    }
    int get_divergence(void* d, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        if (vec4()) LAUNCH((k_divergence_v4<Real>), vec_blocks(g), 256, g, 0, Real(0), Real(0), U(0), U(1), U(2), rhs);
        else LAUNCH((k_divergence<Real>), cell_blocks(), 256, g, 0, Real(0), Real(0), U(0), U(1), U(2), rhs);
        CK(cudaGetLastError());
        if (int rc = allgather_rows(rhs, sizeof(Real))) return rc;
        return grid_out_any(rhs, d, hd);
    }
    int sample_velocity(const void* pos, int64_t n, void* out, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        if (n <= 0) return 0;
        const size_t es = hd == 64 ? 8 : 4;
        // stage: [host copy in | Real in | Real out | host out]
        const size_t a0 = (size_t)n * 3 * es, a1 = (size_t)n * 3 * sizeof(Real);
        if (int rc = need_stage(2 * a0 + 2 * a1)) return rc;
        const void* dp;
        if (int rc = to_device(pos, a0, 0, &dp)) return rc;
        Real* rin = (Real*)((char*)stage + a0);
        Real* rout = (Real*)((char*)stage + a0 + a1);
        void* hout = (char*)stage + a0 + 2 * a1;
        // reuse the generic converters through a flat "grid" of n*3 elements
        if (hd == 64) LAUNCH((k_convert<double, Real>), nblk(n * 3), 256, n * 3, (const double*)dp, rin);
        else LAUNCH((k_convert<float, Real>), nblk(n * 3), 256, n * 3, (const float*)dp, rin);
        if (P.arith == INK_ARITH_FAST) LAUNCH((k_sample<Real, true>), nblk(n), 256, g, C, (long long)n, rin, rout, U(0), U(1), U(2));
        else LAUNCH((k_sample<Real, false>), nblk(n), 256, g, C, (long long)n, rin, rout, U(0), U(1), U(2));
        const bool dout = is_device_ptr(out);
        void* o = dout ? out : hout;
        if (hd == 64) LAUNCH((k_convert<Real, double>), nblk(n * 3), 256, n * 3, rout, (double*)o);
        else LAUNCH((k_convert<Real, float>), nblk(n * 3), 256, n * 3, rout, (float*)o);
        CK(cudaGetLastError());
        if (!dout) CK(cudaMemcpyAsync(out, o, a0, cudaMemcpyDeviceToHost, stream));
        CK(cudaStreamSynchronize(stream));
        return 0;
    }

    // ---- phases --------------------------------------------------------------------------------------
    int read_max_speed(float* out) {
        if (!maxv_valid) {
            CK(cudaMemsetAsync(d_scal, 0, sizeof(unsigned int), stream));
            const unsigned int b = std::min<unsigned int>(cell_blocks(), kSMs * 8);
            LAUNCH((k_max_speed<Real>), b, 256, g, U(0), U(1), U(2), d_scal);
            CK(cudaGetLastError());
            if (int rc = allreduce(d_scal, 1, ncclUint32, ncclMax)) return rc;   // non-negative floats order like their bits
            maxv_valid = true;
        }
        read_small(h_scal, d_scal, sizeof(unsigned int));
        CK(cudaStreamSynchronize(stream));
        float m;
        memcpy(&m, h_scal, 4);
        *out = m;
        return 0;
    }
    float last_grid_maxv = 0.f;
    // System::calcTimeStep (src/waterSolver.cpp:13-38; Q12) — the scalar logic runs on the host in the reference's float arithmetic
    int calc_dt(double ttnr, double* dt_out) override {
        float maxVelocity = 0.f;
        if (int rc = read_max_speed(&maxVelocity)) return rc;
        last_grid_maxv = maxVelocity;
        if (have_phantom) {  // phantom map cells of out-of-grid particles also enter the maximum (Q8)
            const float s = (float)std::sqrt(phantom_v[0] * phantom_v[0] + (phantom_v[1] * phantom_v[1] + phantom_v[2] * phantom_v[2]));
            if (maxVelocity < s) maxVelocity = s;
        }
        double dt;
        if (std::isinf(maxVelocity) || maxVelocity == 0.f) dt = P.dt_min;
        else {
            float timeStep = P.k_cfl * (P.cell_dim / maxVelocity);
            timeStep = std::max(std::min(timeStep, P.dt_max), P.dt_min);
            // the reference assigns the min to its FLOAT local (src/waterSolver.cpp:35): narrowed even when ttnr wins
            timeStep = (float)std::min(ttnr, (double)timeStep);
            dt = timeStep;
        }
        stats.dt = dt;
        if (dt_out) *dt_out = dt;
        return 0;
    }
    void swap_vel() { cur ^= 1; }

    int advect_impl(Real dt, bool allow_skip, int ext = 0) {
        stats.advect_skipped = 0;
        if (allow_skip) {
            // With CFL-limited dt the back-traced cell centre cannot leave its cell, so the nearest-cell copy is the
            // identity: every sampled component is a convex combination of grid values, |v_c| <= max|v|.
            float mv;
            if (int rc = read_max_speed(&mv)) return rc;
            const double h = P.cell_dim, margin = std::min(0.5 * h, 1.0 - 0.5 * h);
            if (margin > 0 && std::isfinite(mv) && std::fabs((double)dt) * (double)mv <= 0.9 * margin) { stats.advect_skipped = 1; return 0; }
        }
        { const Grid q = slab(ext); LAUNCH((k_advect<Real>), cell_blocks(q), 256, q, C, dt, U(0), U(1), U(2), Un(0), Un(1), Un(2)); }
        CK(cudaGetLastError());
        swap_vel();
        maxv_valid = false;
        return 0;
    }
    int advect(double dt) override {
        if (int rc = advect_impl((Real)(float)dt, true)) return rc;
        return stats.advect_skipped ? 0 : allgather_velocity();
    }

    int forces_impl(Real dt, int ext = 0) {
        if (int rc = ensure_seed()) return rc;
        const Real gx = dt * C.g[0], gy = dt * C.g[1], gz = dt * C.g[2];
        { const Grid q = slab(ext);
          if (vec4() && C.R <= 4) LAUNCH((k_dilate_add_gravity_v4<Real>), vec_blocks(q), 256, q, C.R, seed, mark, gx, gy, gz, U(0), U(1), U(2));
          else LAUNCH((k_dilate_add_gravity<Real>), cell_blocks(q), 256, q, C.R, seed, mark, gx, gy, gz, U(0), U(1), U(2)); }
        CK(cudaGetLastError());
        if (oob_now) { phantom_v[0] += gx; phantom_v[1] += gy; phantom_v[2] += gz; }
        maxv_valid = false;
        return 0;
    }
    int forces(double dt) override {
        if (int rc = forces_impl((Real)(float)dt)) return rc;
        return allgather_velocity();
    }

    int viscosity_impl(Real dt, int ext = 0) {
        const Real s = dt * C.visc;
        { const Grid q = slab(ext);
          if (vec4()) LAUNCH((k_viscosity_v4<Real>), vec_blocks(q), 256, q, s, U(0), U(1), U(2), Un(0), Un(1), Un(2));
          else LAUNCH((k_viscosity<Real>), cell_blocks(q), 256, q, s, U(0), U(1), U(2), Un(0), Un(1), Un(2)); }
        CK(cudaGetLastError());
        swap_vel();
        maxv_valid = false;
        return 0;
    }
    int viscosity(double dt) override {
        if (int rc = viscosity_impl((Real)(float)dt)) return rc;
        return allgather_velocity();
    }

    int pressure_rhs(double dt_, void* b, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        const Real dt = (Real)(float)dt_;
        const Real scale = (C.rho * C.h) / dt;
        if (vec4()) LAUNCH((k_divergence_v4<Real>), vec_blocks(g), 256, g, 1, scale, C.patm, U(0), U(1), U(2), rhs);
        else LAUNCH((k_divergence<Real>), cell_blocks(), 256, g, 1, scale, C.patm, U(0), U(1), U(2), rhs);
        CK(cudaGetLastError());
        if (int rc = allgather_rows(rhs, sizeof(Real))) return rc;
        return grid_out_any(rhs, b, hd);
    }

    // ---- multigrid hierarchy ------------------------------------------------------------------------
    int mg_setup() {
        if (const char* e = getenv("INK_TRACE_ITER")) trace_iter = atoi(e);
        if (const char* e = getenv("INK_HOSTTIME")) host_timing = atoi(e) != 0;
        if (const char* e = getenv("INK_NO_GRAPH")) use_graph = (atoi(e) == 0);
        if (const char* e = getenv("INK_MG3_FUSED_DOWN")) mg_fused_down = atoi(e) != 0;
        if (const char* e = getenv("INK_MG3_COARSEST")) mg_coarsest = atoi(e) != 0;
        if (const char* e = getenv("INK_MG3_UP_TILE")) mg_up_tile = atoi(e);
        if (const char* e = getenv("INK_MG3_DOWN_TILE")) mg_down_tile = atoi(e);
        if (const char* e = getenv("INK_MG_NODONE")) mg_nodone = atoi(e) != 0;
        CK(cudaFuncSetAttribute((k_mg3_up_tile<Real, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mg3_up_tile_smem<Real>()));
        CK(cudaFuncSetAttribute((k_mg3_up_tile<Real, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)mg3_up_tile_smem<Real>()));
        if (const char* e = getenv("INK_MG3_UNFUSED")) mg_fuse = (atoi(e) == 0);
        if (const char* e = getenv("INK_MG_TAIL_KB")) mg_tail_cap = (size_t)std::max(1, atoi(e)) * 1024;
        // measured (B200, 128x256x128): a tail holding only the coarsest (<= 1024-cell) level beats a deeper single-block tail
        // (2.59 vs 2.73 ms/step): one block of 1024 threads is issue-bound on the 8192-cell level
        if (mg_tri && !getenv("INK_MG_TAIL_KB")) mg_tail_cap = 32 * 1024;
        const size_t smem_cap = std::min<size_t>(mg_tail_cap, 200 * 1024);
        std::vector<MgDims> abs_d;
        je0 = g.j0; je1 = g.j1; dot0 = g.j0; dot1 = g.j1;
        if (world > 1) {
            int min_rows = g.Y;
            for (int r = 0; r < world; ++r) min_rows = std::min(min_rows, slab_lo[r + 1] - slab_lo[r]);
            int ovl_max = 8;   // INK_SCHWARZ_OVL: overlap rows of the Schwarz preconditioner (always <= half the thinnest slab)
            if (const char* e = getenv("INK_SCHWARZ_OVL")) ovl_max = std::max(1, atoi(e));
            ovl = std::max(1, std::min(ovl_max, min_rows / 2));
            if (rank > 0) { je0 = g.j0 - ovl; dot0 = g.j0 + ovl; }
            if (rank + 1 < world) { je1 = g.j1 + ovl; dot1 = g.j1 - ovl; }
            // sqrt of the linear partition of unity across each 2*ovl-row band (two ranks' weights sum to 1 on every row)
            std::vector<Real> sw((size_t)g.Y, Real(0));
            for (int j = je0; j < je1; ++j) {
                double w = 1.0;
                if (rank > 0 && j < g.j0 + ovl) w = std::min(w, (j - (g.j0 - ovl) + 0.5) / (2.0 * ovl));
                if (rank + 1 < world && j >= g.j1 - ovl) w = std::min(w, ((g.j1 + ovl) - j - 0.5) / (2.0 * ovl));
                sw[j] = (Real)std::sqrt(w);
            }
            CK(cudaMalloc(&d_sw, (size_t)g.Y * sizeof(Real)));
            CK(cudaMemcpyAsync(d_sw, sw.data(), (size_t)g.Y * sizeof(Real), cudaMemcpyHostToDevice, stream));
            CK(cudaStreamSynchronize(stream));   // sw is a local
            const size_t band = (size_t)2 * ovl * g.sy * sizeof(Real);
            for (int b = 0; b < 2; ++b) { CK(cudaMalloc(&band_save[b], band)); CK(cudaMalloc(&band_recv[b], band)); }
        }
        abs_d.push_back(MgDims{g.X, g.Y, g.Z, je0, je1});
        auto cells_local = [](const MgDims& d) { return (long long)d.X * (d.j1 - d.j0) * d.Z; };
        // tail = local-coordinate levels below (and including) some level; find the first level whose tail fits in shared memory
        auto build_tail = [&](const MgDims& top, MgTailArgs& T) -> size_t {
            MgDims d{top.X, top.j1 - top.j0, top.Z, 0, top.j1 - top.j0};
            T.nlev = 0;
            long long total = 0;
            while (true) {
                T.d[T.nlev++] = d;
                total += 3ll * d.X * d.Y * d.Z;
                if ((long long)d.X * d.Y * d.Z <= (mg_tri ? kMg3CoarsestCells : 16) || T.nlev == kMgMaxLevels) break;
                d = MgDims{(d.X + 1) / 2, (d.Y + 1) / 2, (d.Z + 1) / 2, 0, (d.Y + 1) / 2};
            }
            return (size_t)total * sizeof(Real);
        };
        while (true) {
            mg_tail_smem = build_tail(abs_d.back(), mg_tail);
            if (mg_tail_smem <= smem_cap) break;
            if ((int)abs_d.size() >= kMgMaxLevels) { err = "multigrid: grid too large for the level budget"; return 2; }
            const MgDims& d = abs_d.back();
            abs_d.push_back(MgDims{(d.X + 1) / 2, (d.Y + 1) / 2, (d.Z + 1) / 2, d.j0 >> 1, (d.j1 + 1) >> 1});
            (void)cells_local;
        }
        mg_lt = (int)abs_d.size() - 1;
        mg_tail.s0 = (float)(1 << mg_lt);
        mg_tail.final_dot = (mg_lt == 0) ? 1 : 0;
        mg_tail.dj0 = dot0 - je0; mg_tail.dj1 = dot1 - je0;
        lv.resize(abs_d.size());
        for (size_t l = 0; l < abs_d.size(); ++l) {
            lv[l].d = abs_d[l];
            lv[l].s = (Real)(1 << l);
            const size_t nb = (size_t)abs_d[l].X * abs_d[l].Y * abs_d[l].Z * sizeof(Real);
            Real** slots[3] = {&lv[l].xa, &lv[l].xb, &lv[l].r};
            for (int a = 0; a < 3; ++a) {
                *slots[a] = nullptr;
                if (a == 2 && l == 0) continue;           // level-0 rhs is the CG residual itself
                if (a == 1 && (int)l == mg_lt) continue;  // the tail's top level only needs its result buffer
                CK(cudaMalloc(slots[a], nb));
                CK(cudaMemsetAsync(*slots[a], 0, nb, stream));
                mg_alloc.push_back(*slots[a]);
            }
        }
        // The attribute belongs to the FUNCTION, not to this context object: several Systems of different sizes may live in one
        // process (tests/mgpu_check.py: a distributed and a single-GPU one), and a later, smaller one must not lower the limit
        // under an earlier one's launches ("invalid argument").  Keep a process-wide running maximum, never below the default.
        static std::atomic<size_t> tail_attr{48 * 1024};
        size_t want = tail_attr.load();
        while (want < mg_tail_smem && !tail_attr.compare_exchange_weak(want, mg_tail_smem)) {}
        want = std::max(want, mg_tail_smem);
        CK(cudaFuncSetAttribute(k_mg_tail<Real>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)want));
        CK(cudaFuncSetAttribute(k_mg3_tail<Real>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)want));
        // Coarse levels (<= kMg3CoarseCells cells: launch-latency bound, ~4.5 us per kernel for < 1 us of work) in ONE launch.
        // INK_MG_COARSE=2: one thread-block cluster of 16 CTAs with the hardware cluster barrier between passes (falls back to 8
        // CTAs, then to per-level kernels, if the device refuses the cluster).  INK_MG_COARSE=1: the round-1 variant, a 32-block
        // co-resident grid with a global-memory barrier.  Default 0: per-level kernels.  Measured on B200 at 128x256x128 (round 2,
        // profiles/r2_c_*): the PDL-chained launch train that uses all 148 SMs (pressure phase 0.97 ms) beats both single-launch
        // variants (cluster of 16: 1.29 ms, of 8: 1.36 ms; grid barrier: slower still) — on 16 SMs the passes themselves are
        // latency-bound (a few warps per scheduler walking L2-latency load chains), which costs more than the launches save.
        int coarse_mode = 0;
        if (const char* e = getenv("INK_MG_COARSE")) coarse_mode = mg_tri ? atoi(e) : 0;
        int coarse_cells = kMg3CoarseCells;
        if (const char* e = getenv("INK_MG_COARSE_CELLS")) coarse_cells = std::max(1, atoi(e));
        if (coarse_mode != 0) {
            for (int l = 1; l < mg_lt; ++l) {
                const MgDims& d = lv[l].d;
                if ((long long)d.X * (d.j1 - d.j0) * d.Z <= coarse_cells) { mg_lc = l; break; }
            }
            if (mg_lc >= 1 && mg_lt - mg_lc + 1 > kMg3CoarseMaxLevels) mg_lc = -1;
        }
        if (mg_lc >= 1 && coarse_mode == 1) {
            CK(cudaMalloc(&mg_bar, 2 * sizeof(unsigned int)));
            CK(cudaMemsetAsync(mg_bar, 0, 2 * sizeof(unsigned int), stream));
            CK(cudaFuncSetAttribute((k_mg3_coarse<Real, false>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)want));
        } else if (mg_lc >= 1) {
            CK(cudaFuncSetAttribute((k_mg3_coarse<Real, true>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)want));
            mg_cluster = kMg3ClusterCtas;
            if (const char* e = getenv("INK_MG_CLUSTER_CTAS")) mg_cluster = std::max(1, std::min(16, atoi(e)));
            if (mg_cluster > 8 && cudaFuncSetAttribute((k_mg3_coarse<Real, true>), cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
                cudaGetLastError();
                mg_cluster = 8;
            }
            // can the device co-schedule such a cluster at all?
            while (mg_cluster >= 2) {
                cudaLaunchConfig_t cfg = {};
                cfg.gridDim = dim3(mg_cluster); cfg.blockDim = dim3(kMg3CoarseThreads); cfg.dynamicSmemBytes = mg_tail_smem;
                cudaLaunchAttribute at[1];
                at[0].id = cudaLaunchAttributeClusterDimension;
                at[0].val.clusterDim.x = mg_cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                cfg.attrs = at; cfg.numAttrs = 1;
                int ncl = 0;
                if (cudaOccupancyMaxActiveClusters(&ncl, (k_mg3_coarse<Real, true>), &cfg) == cudaSuccess && ncl >= 1) break;
                cudaGetLastError();
                mg_cluster /= 2;
            }
            if (mg_cluster < 2) { mg_cluster = 0; mg_lc = -1; }
        }
        use_mg = true;
        return 0;
    }
    static unsigned int mg_blocks(const MgDims& d) {
        const long long n = (long long)d.X * (d.j1 - d.j0) * d.Z;
        return (unsigned int)std::min<long long>(kMgBlocks, std::max<long long>(1, (n + 255) / 256));
    }
    // z = B r : one V-cycle; the result is lv[0].xa; the last kernel also sets S->rz_old, S->rz = <r, z>
    int mg_vcycle(const Real* r0) {
        auto rhs = [&](int l) -> const Real* { return l == 0 ? r0 : lv[l].r; };
        if (mg_tri) {
            // INK_MG_NODONE=1: the V-cycle kernels that touch S only for the `done` early exit get S = nullptr and skip that dependent
            // load (~0.5 us in front of every one of ~11 short kernels per iteration); after convergence they then recompute z from
            // an unchanged r instead of exiting — harmless (x, r and the scalars are guarded by the CG kernels and the FINAL up kernel)
            const CgScal* Sd = mg_nodone ? nullptr : S;
            // V(1,1), trilinear transfer: 3 launches per global level (residual, restrict | up), xb = residual scratch
            const int l_split = mg_lc >= 1 ? mg_lc : mg_lt;   // levels below l_split: one kernel per pass
            for (int l = 0; l < l_split; ++l) {
                const MgDims &df = lv[l].d, &dc = lv[l + 1].d;
                if (mg_down_tile && df.Z % 4 == 0 && (long long)df.X * df.Y * df.Z < (1ll << 31) && (l >= 1 || mg_down_tile == 2)) {
                    // residual + restriction through a shared-memory tile, one launch (mg_tri.cuh: k_mg3_down_tile)
                    const int ntx = (dc.X + kDnCX - 1) / kDnCX, nty = (dc.j1 - dc.j0 + kDnCY - 1) / kDnCY, ntz = (dc.Z + kDnCZ - 1) / kDnCZ;
                    const unsigned int nbt = (unsigned int)std::min<long long>((long long)ntx * nty * ntz, 148 * 32);
                    LAUNCH((k_mg3_down_tile<Real>), nbt, 256, Sd, df, dc, rhs(l), lv[l + 1].r, ntx, nty, ntz);
                    TR(l == 0 ? "down L0" : l == 1 ? "down L1" : "down L2+");
                    continue;
                }
                if (l >= 1 && mg_fused_down) {    // below the finest level a kernel costs its launch, not its traffic: one launch per level
                    LAUNCH((k_mg3_down_fused<Real>), mg_blocks(dc), 256, Sd, df, dc, rhs(l), lv[l + 1].r);
                    TR(l == 1 ? "down L1" : "down L2+");
                    continue;
                }
                if (df.Z % 4 == 0 && mg_fuse) {   // residual + z-restriction fused (xb holds the half-size z-restricted residual)
                    const long long nv = (long long)df.X * (df.j1 - df.j0) * (df.Z / 4);
                    const unsigned int nbr = (unsigned int)std::min<long long>(kMgBlocks, std::max<long long>(1, (nv + 255) / 256));
                    if (df.Z / 4 > 32) LAUNCH((k_mg3_residual_rz_v4_blk<Real>), nbr, 256, Sd, df, rhs(l), lv[l].xb);   // rows longer than a warp of quads
                    else LAUNCH((k_mg3_residual_rz_v4<Real>), nbr, 256, Sd, df, rhs(l), lv[l].xb);
                    if (dc.Z % 4 == 0) {
                        const long long nc4 = (long long)dc.X * (dc.j1 - dc.j0) * (dc.Z / 4);
                        LAUNCH((k_mg3_restrict_xy_v4<Real>), (unsigned int)std::min<long long>(kMgBlocks, std::max<long long>(1, (nc4 + 255) / 256)), 256, Sd, df, dc, lv[l].xb, lv[l + 1].r);
                    } else LAUNCH((k_mg3_restrict_xy<Real>), mg_blocks(dc), 256, Sd, df, dc, lv[l].xb, lv[l + 1].r);
                    TR(l == 0 ? "down L0" : l == 1 ? "down L1" : "down L2+");
                    continue;
                }
                if (df.Z % 4 == 0) {
                    const long long nv = (long long)df.X * (df.j1 - df.j0) * (df.Z / 4);
                    LAUNCH((k_mg3_residual_v4<Real>), (unsigned int)std::min<long long>(kMgBlocks, std::max<long long>(1, (nv + 255) / 256)), 256, Sd, df, rhs(l), lv[l].xb);
                } else LAUNCH((k_mg3_residual<Real>), mg_blocks(df), 256, Sd, df, rhs(l), lv[l].xb);
                LAUNCH((k_mg3_restrict<Real>), mg_blocks(dc), 256, Sd, df, dc, lv[l].xb, lv[l + 1].r);
            }
            if (mg_lc >= 1) {   // levels mg_lc .. mg_lt (down, tail, up) in one persistent launch
                Mg3CoarseArgs<Real> A{};
                A.nlev = mg_lt - mg_lc + 1;
                for (int q = 0; q < A.nlev; ++q) A.lv[q] = lv[mg_lc + q];
                A.tail = mg_tail;
                A.tail_off = (long long)lv[mg_lt].d.j0 * lv[mg_lt].d.X * lv[mg_lt].d.Z;
                A.bar = mg_bar;
                if (mg_cluster > 0) launch_cluster(k_mg3_coarse<Real, true>, mg_cluster, kMg3CoarseThreads, mg_tail_smem, stream, S, A);
                else launch_k(k_mg3_coarse<Real, false>, kMg3CoarseBlocks, kMg3CoarseThreads, mg_tail_smem, stream, S, A);
                ++stats.kernel_launches;
                TR("coarse levels");
            } else {
                const MgDims& d = lv[mg_lt].d;
                const long long off = (long long)d.j0 * d.X * d.Z;
                if (mg_coarsest && mg_tail.nlev == 1 && !mg_tail.final_dot && (long long)d.X * (d.j1 - d.j0) * d.Z <= kMg3CoarsestCells)
                    launch_k(k_mg3_coarsest<Real>, 1, kMg3CoarsestThreads, 0, stream, Sd, d, lv[mg_lt].s, rhs(mg_lt) + off, lv[mg_lt].xa + off);
                else launch_k(k_mg3_tail<Real>, 1, kMgTailThreads, mg_tail_smem, stream, S, mg_tail, rhs(mg_lt) + off, lv[mg_lt].xa + off);
                ++stats.kernel_launches;
                TR("tail");
            }
            for (int l = l_split - 1; l >= 0; --l) {
                const MgDims &df = lv[l].d, &dc = lv[l + 1].d;
                // one coarse cell per thread, no grid-stride tail (a thread's work is ~700 instructions); the fp64 partials of
                // block_sum_last cap the FINAL launch at kCgMaxBlocks blocks
                const long long ncc = (long long)dc.X * (dc.j1 - dc.j0) * dc.Z;
                const unsigned int ub = (unsigned int)std::max<long long>(1, (ncc + 255) / 256);
                // levels with enough tiles to fill the GPU: the shared-memory tile version (38 instead of 85 instructions per fine cell)
                const int ntx = (df.X + kUpTX - 1) / kUpTX, nty = (df.j1 - (df.j0 & ~1) + kUpTY - 1) / kUpTY, ntz = (df.Z + kUpTZ - 1) / kUpTZ;
                const long long ntile = (long long)ntx * nty * ntz;
                if (mg_up_tile && df.Z % 4 == 0 && (long long)df.X * df.Y * df.Z < (1ll << 31) && ntile >= (mg_up_tile == 2 ? 1 : 2 * kSMs)) {
                    const unsigned int nb = (unsigned int)std::min<long long>(ntile, kCgMaxBlocks);
                    if (l == 0) launch_k((k_mg3_up_tile<Real, true>), nb, 256, mg3_up_tile_smem<Real>(), stream, S, df, dc, lv[l].s, rhs(l), (const Real*)lv[l + 1].xa, lv[l].xa, dot0, dot1, ntx, nty, ntz);
                    else launch_k((k_mg3_up_tile<Real, false>), nb, 256, mg3_up_tile_smem<Real>(), stream, (CgScal*)Sd, df, dc, lv[l].s, rhs(l), (const Real*)lv[l + 1].xa, lv[l].xa, 0, 0, ntx, nty, ntz);
                    ++stats.kernel_launches;
                }
                else if (l == 0) LAUNCH((k_mg3_up<Real, true>), std::min<unsigned int>(ub, kCgMaxBlocks), 256, S, df, dc, lv[l].s, rhs(l), lv[l + 1].xa, lv[l].xa, dot0, dot1);
                else LAUNCH((k_mg3_up<Real, false>), ub, 256, (CgScal*)Sd, df, dc, lv[l].s, rhs(l), lv[l + 1].xa, lv[l].xa, 0, 0);
                TR(l == 0 ? "up L0" : l == 1 ? "up L1" : "up L2+");
            }
            CK(cudaGetLastError());
            return 0;
        }
        auto vblocks = [](const MgDims& d, int per) {   // one thread per `per` cells
            const long long n = (long long)d.X * (d.j1 - d.j0) * (d.Z / per);
            return (unsigned int)std::min<long long>(kMgBlocks, std::max<long long>(1, (n + 255) / 256));
        };
        for (int l = 0; l < mg_lt; ++l) {
            const MgDims &df = lv[l].d, &dc = lv[l + 1].d;
            if (df.Z % 4 == 0) {
                LAUNCH((k_mg_presmooth_v4<Real>), vblocks(df, 4), 256, S, df, lv[l].s, rhs(l), lv[l].xa);
                LAUNCH((k_mg_residual_restrict_v4<Real>), vblocks(dc, 2), 256, S, df, dc, lv[l].s, rhs(l), lv[l].xa, lv[l + 1].r);
            } else {
                LAUNCH((k_mg_presmooth<Real>), mg_blocks(df), 256, S, df, lv[l].s, rhs(l), lv[l].xa);
                LAUNCH((k_mg_residual_restrict<Real>), mg_blocks(dc), 256, S, df, dc, lv[l].s, rhs(l), lv[l].xa, lv[l + 1].r);
            }
        }
        {
            const MgDims& d = lv[mg_lt].d;
            const long long off = (long long)d.j0 * d.X * d.Z;
            launch_k(k_mg_tail<Real>, 1, kMgTailThreads, mg_tail_smem, stream, S, mg_tail, rhs(mg_lt) + off, lv[mg_lt].xa + off);
            ++stats.kernel_launches;
        }
        for (int l = mg_lt - 1; l >= 0; --l) {
            const MgDims &df = lv[l].d, &dc = lv[l + 1].d;
            if (df.Z % 4 == 0) {
                LAUNCH((k_mg_prolong_smooth_v4<Real>), vblocks(df, 4), 256, S, df, dc, lv[l].s, rhs(l), lv[l].xa, lv[l + 1].xa, lv[l].xb);
                if (l == 0) LAUNCH((k_mg_smooth_v4<Real, true>), vblocks(df, 4), 256, S, df, lv[l].s, rhs(l), lv[l].xb, lv[l].xa, dot0, dot1);
                else LAUNCH((k_mg_smooth_v4<Real, false>), vblocks(df, 4), 256, S, df, lv[l].s, rhs(l), lv[l].xb, lv[l].xa, 0, 0);
            } else {
                LAUNCH((k_mg_prolong_smooth<Real>), mg_blocks(df), 256, S, df, dc, lv[l].s, rhs(l), lv[l].xa, lv[l + 1].xa, lv[l].xb);
                if (l == 0) LAUNCH((k_mg_smooth<Real, true>), mg_blocks(df), 256, S, df, lv[l].s, rhs(l), lv[l].xb, lv[l].xa, dot0, dot1);
                else LAUNCH((k_mg_smooth<Real, false>), mg_blocks(df), 256, S, df, lv[l].s, rhs(l), lv[l].xb, lv[l].xa, 0, 0);
            }
        }
        CK(cudaGetLastError());
        return 0;
    }

    // z = B r (result lv[0].xa, valid on own rows and `ovl` rows into each neighbour).  One GPU: one V-cycle.  Multi-GPU:
    // symmetric weighted overlapping Schwarz around the slab-local V-cycle — two band exchanges, no traffic inside the cycle.
    int precondition(Real* r, const ArFold& fold = ArFold{}) {
        if (world == 1) { const int rc = mg_vcycle(r); TR("vcycle"); return rc; }
        const long long sy = g.sy;
        const size_t brow = (size_t)ovl * sy;
        // r on the `ovl` rows beyond each interface: my bottom rows go to rank-1, my top rows to rank+1
        // (peer-memory path: the producer of r already pushed those rows into the neighbours' arenas, make_push())
        if (!p2p)
            if (int rc = halo_exchange(r, ovl, ovl)) return rc;
        TR("r halo");
        // 16-byte accesses when a grid row is a multiple of 16 bytes (always true on the peer-memory path)
        constexpr int kVec = 16 / (int)sizeof(Real);
        const bool bvec = ((size_t)sy * sizeof(Real)) % 16 == 0;
#define LAUNCH_BAND(kern, ...) do { if (bvec) LAUNCH((kern<Real, kVec>), __VA_ARGS__); else LAUNCH((kern<Real, 1>), __VA_ARGS__); } while (0)
        struct Band { int ja, jb, peer; };
        Band bands[2] = {{g.j0 - ovl, g.j0 + ovl, rank - 1}, {g.j1 - ovl, g.j1 + ovl, rank + 1}};
        const bool has[2] = {rank > 0, rank + 1 < world};
        // band kernels: few blocks per side (each ends with a system-scope fence when it pushes to a peer), grid-stride; ONE launch
        // per stage covers both sides of an interior rank (blocks [0, nb) side 0, [nb, 2 nb) side 1)
        const unsigned int bb = (unsigned int)std::min<long long>(2 * kSMs, (2 * (long long)brow + 255) / 256);
        const unsigned long long z_seq = p2p ? ++p2p_seq[1] : 0ull;
        BandArgs<Real> A{};
        A.nb = (int)bb; A.r_seq = r_seq; A.z_seq = z_seq;
        for (int b = 0; b < 2; ++b)
            if (has[b]) {
                BandSide<Real>& B = A.s[A.nsides++];
                B.ja = bands[b].ja; B.jb = bands[b].jb; B.save = band_save[b];
                if (p2p) {
                    B.halo_src = reinterpret_cast<const Real*>(arena + off_halo[b]); B.halo_j0 = (b == 0) ? g.j0 - ovl : g.j1;
                    B.wait_r = &head_of(rank)->flag[0][b];
                    B.push_dst = reinterpret_cast<Real*>(peers.base[bands[b].peer] + off_band[1 - b]);
                    B.push_flag = &head_of(bands[b].peer)->flag[1][1 - b];
                    B.ticket = &head_of(rank)->ticket[2 + b];
                    B.recv = reinterpret_cast<const Real*>(arena + off_band[b]);   // filled by the neighbour's k_band_restore_weight
                    B.wait_z = &head_of(rank)->flag[1][b];
                } else B.recv = band_recv[b];
            }
        const unsigned int grid = bb * (unsigned int)A.nsides;
        LAUNCH_BAND(k_band_scale, grid, 256, S, sy, A, g.j0, g.j1, d_sw, r);
        TR("band scale");
        if (int rc = mg_vcycle(r)) return rc;
        TR("vcycle");
        Real* z = lv[0].xa;
        LAUNCH_BAND(k_band_restore_weight, grid, 256, S, sy, A, g.j0, g.j1, d_sw, r, z);
        CK(cudaGetLastError());
        TR("band weight");
        if (!p2p) {
            NK(nccl().GroupStart());
            for (int b = 0; b < 2; ++b)
                if (has[b]) {
                    NK(nccl().Send(z + (size_t)bands[b].ja * sy, 2 * brow, nc_real(), bands[b].peer, comm, stream));
                    NK(nccl().Recv(band_recv[b], 2 * brow, nc_real(), bands[b].peer, comm, stream));
                }
            NK(nccl().GroupEnd());
        }
        LAUNCH_BAND(k_band_combine, grid, 256, S, sy, A, g.j0, g.j1, (const Real*)r, z, fold);   // + the folded all-reduce of <r,r>, <r,z>
#undef LAUNCH_BAND
        TR("z band combine");
        CK(cudaGetLastError());
        return 0;
    }

    // solve M x = f (M = -A) by (preconditioned) CG; x holds the initial guess.
    // Multi-GPU: every rank iterates on its own rows; per iteration 2 all-reduces (<d,w>; <r,r> & <r,z> together) and one
    // boundary-row exchange of z.  The preconditioner is slab-local (no traffic).
    int cg_post(int stage) {
        launch_k(k_cg_post, 1, 32, 0, stream, S, stage, use_mg ? 1 : 0);
        ++stats.kernel_launches;
        return 0;
    }
    // capture `body(dc)` once per parity into a graph (thread-local capture mode: other host threads of the process are not
    // affected), then replay it.  The cached pair belongs to one solution vector x; another x recaptures.
    // A capture or an instantiation the runtime refuses (a driver without programmatic-dependency edges in captures, another
    // capture open on this thread, ...) is not an error of the solve: the ctx drops to plain launches for good (use_graph = false,
    // one line under INK_VERBOSE) and runs the iteration directly.  Only what the body itself returns is a hard failure.
    template <class Body>
    int graph_fallback(int dc, Body& body, const char* what, cudaError_t e) {
        cudaGetLastError();   // clear the sticky-free error state the failed call left
        use_graph = false;
        for (auto& ge : it_graph) if (ge) { cudaGraphExecDestroy(ge); ge = nullptr; }
        if (getenv("INK_VERBOSE")) fprintf(stderr, "[inkb200] %s failed (%s): PCG iterations run as plain launches from here on\n", what, cudaGetErrorString(e));
        return body(dc);
    }
    template <class Body>
    int replay_iteration(int dc, const void* x, Body& body) {
        if (!use_graph) return body(dc);   // fell back earlier in this solve
        if (it_graph_x != x) {
            for (auto& ge : it_graph) if (ge) { cudaGraphExecDestroy(ge); ge = nullptr; }
            it_graph_x = x;
        }
        if (!it_graph[dc]) {
            const long long before = stats.kernel_launches;
            cudaGraph_t gr = nullptr;
            const cudaError_t be = g_fail_capture ? cudaErrorStreamCaptureUnsupported : cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal);
            if (be != cudaSuccess) return graph_fallback(dc, body, "cudaStreamBeginCapture", be);
            const int rc = body(dc);
            const cudaError_t ce = cudaStreamEndCapture(stream, &gr);
            const long long nodes = stats.kernel_launches - before;
            stats.kernel_launches = before;   // capturing launched nothing
            if (rc) { if (gr) cudaGraphDestroy(gr); cudaGetLastError(); return rc; }
            if (ce != cudaSuccess || !gr) return graph_fallback(dc, body, "graph capture of the PCG iteration", ce);
            const cudaError_t ie = cudaGraphInstantiate(&it_graph[dc], gr, 0);
            cudaGraphDestroy(gr);
            if (ie != cudaSuccess) { it_graph[dc] = nullptr; return graph_fallback(dc, body, "cudaGraphInstantiate", ie); }
            it_graph_nodes = nodes;
        }
        CK(cudaGraphLaunch(it_graph[dc], stream));
        stats.kernel_launches += it_graph_nodes;
        ++stats.graph_launches;
        return 0;
    }
    int cg_solve(const Real* f, Real* x) {
        const unsigned int nb = cg_blocks();
        const bool dist = world > 1;
        const int hlo = (dist && rank > 0) ? 1 : 0, hhi = (dist && rank + 1 < world) ? 1 : 0;
        if (int rc = halo_exchange(x, 1, 1)) return rc;     // warm start: the initial residual needs x on the neighbour rows
        { const P2pPush push = make_push();
          const ArFold fold = make_fold(0, 4, 0);
          if (vec4()) LAUNCH((k_cg_init<Real, 4>), nb, 256, g, S, f, x, cg_r, push, fold);
          else LAUNCH((k_cg_init<Real, 1>), nb, 256, g, S, f, x, cg_r, push, fold);
          CK(cudaGetLastError());
          if (dist && !fold.on) {
              if (int rc = ar_post(0, 4, 0)) return rc;
          } }
        const size_t head = offsetof(CgScal, partial);
        // Steady state (one GPU, preconditioned, the last solve took > 1 iteration): do NOT wait for the initial residual —
        // the first batch below is launched blind anyway.  (Measured on B200 at 128x256x128: 1.95 vs 1.97 ms per step, i.e.
        // within noise — the GPU was not waiting on this round trip; kept because it is one host sync less per step.)
        // The f == 0 case is then caught at the first poll (the blind iterations ran on bb = 0; x is overwritten with the
        // exact answer 0 all the same).
        const bool blind = use_mg && !dist && prev_iters > 1;
        auto zero_rhs = [&]() -> int {   // f == 0: the solution is exactly 0
            CK(cudaMemsetAsync(x, 0, (size_t)g.n * sizeof(Real), stream));
            stats.cg_iters = 0; stats.cg_rel_residual = 0; stats.cg_converged = 1;
            return 0;
        };
        if (blind) hS->done = 0;   // the mirror still holds the previous solve's head; refreshed at the first poll
        else {
            read_small(hS, S, head);
            CK(cudaStreamSynchronize(stream));
            if (hS->bb == 0.0) return zero_rhs();
        }
        int dcur = 0, launched = 0;
        Real* zsrc = cg_r;
        if (use_mg) {
            const ArFold fold = make_fold(2, 1, 3);
            if (int rc = precondition(cg_r, fold)) return rc;
            zsrc = lv[0].xa;
            if (dist && !fold.on) {
                if (int rc = ar_post(2, 1, 3)) return rc;
            }
        } else if (int rc = halo_exchange(zsrc, 1, 1)) return rc;
        // host polls convergence every `every` iterations; kernels launched after convergence are no-ops (S->done).
        // With the preconditioner the count is small and steady from step to step: launch last solve's count first.
        const int every = use_mg ? 2 : std::max(1, P.cg_check_every);
        // one PCG iteration: direction + matvec, update, preconditioner (+ the distributed reductions / exchanges)
        auto iteration_body = [&](int dc) -> int {
            TR("start");
            const ArFold fold1 = make_fold(0, 1, 1);
            if (mv_tile && vec4() && g.n < (1ll << 31)) {   // d' staged once per cell in shared memory (cg.cuh: k_cg_dir_matvec_tile)
                const int ntx = (g.X + kMvTX - 1) / kMvTX, nty = (g.j1 - g.j0 + kMvTY - 1) / kMvTY, ntz = (g.Z + kMvTZ - 1) / kMvTZ;
                const unsigned int nbt = (unsigned int)std::min<long long>((long long)ntx * nty * ntz, kCgMaxBlocks);
                launch_k((k_cg_dir_matvec_tile<Real>), nbt, 256, cg_matvec_tile_smem<Real>(), stream, g, S, (const Real*)zsrc, (const Real*)cg_d[dc], cg_d[dc ^ 1], cg_w, hlo, hhi, fold1, ntx, nty, ntz);
                ++stats.kernel_launches;
            }
            else if (vec4()) LAUNCH((k_cg_dir_matvec<Real, 4>), nb, 256, g, S, zsrc, cg_d[dc], cg_d[dc ^ 1], cg_w, hlo, hhi, fold1);
            else LAUNCH((k_cg_dir_matvec<Real, 1>), nb, 256, g, S, zsrc, cg_d[dc], cg_d[dc ^ 1], cg_w, hlo, hhi, fold1);
            TR("matvec");
            if (dist && !fold1.on) {
                if (int rc = ar_post(0, 1, 1)) return rc;
            }
            TR("allreduce dw");
            const P2pPush push = make_push();
            if (vec4()) {
                if (use_mg) LAUNCH((k_cg_update<Real, 4, true>), nb, 256, g, S, x, cg_r, cg_d[dc ^ 1], cg_w, push);
                else LAUNCH((k_cg_update<Real, 4, false>), nb, 256, g, S, x, cg_r, cg_d[dc ^ 1], cg_w, push);
            } else {
                if (use_mg) LAUNCH((k_cg_update<Real, 1, true>), nb, 256, g, S, x, cg_r, cg_d[dc ^ 1], cg_w, push);
                else LAUNCH((k_cg_update<Real, 1, false>), nb, 256, g, S, x, cg_r, cg_d[dc ^ 1], cg_w, push);
            }
            TR("update");
            const ArFold fold2 = use_mg ? make_fold(1, 2, 2) : ArFold{};
            if (use_mg)
                if (int rc = precondition(cg_r, fold2)) return rc;
            if (dist) {
                if (!fold2.on)
                    if (int rc = ar_post(1, 2, 2)) return rc;
                TR("allreduce rr,rz");
                if (!use_mg)
                    if (int rc = halo_exchange(zsrc, 1, 1)) return rc;
            }
            return 0;
        };
        // graph replay: single GPU (the distributed body carries host-side sequence numbers in its kernel arguments), not while tracing
        const bool graph_ok = use_graph && !dist && trace_iter < 0;
        while (!hS->done && launched < P.cg_max_iters) {
            int batch = std::min(every, P.cg_max_iters - launched);
            const auto t_batch = std::chrono::steady_clock::now();
            if (use_mg && launched == 0 && prev_iters > 1) batch = std::min(prev_iters, P.cg_max_iters);
            for (int it = 0; it < batch; ++it) {
                tracing = (trace_iter >= 0 && launched + it == trace_iter);
                if (graph_ok) {
                    if (int rc = replay_iteration(dcur, x, iteration_body)) return rc;
                } else if (int rc = iteration_body(dcur)) return rc;
                dcur ^= 1;
                tracing = false;
            }
            launched += batch;
            CK(cudaGetLastError());
            read_small(hS, S, head);
            const auto t_enq = std::chrono::steady_clock::now();
            CK(cudaStreamSynchronize(stream));
            if (host_timing && rank == 0) {   // INK_HOSTTIME=1: is the host (launch rate) or the device the bottleneck of the solve?
                const auto t_done = std::chrono::steady_clock::now();
                fprintf(stderr, "[inkb200 host] CG batch of %d iterations: enqueue %.0f us, then blocked %.0f us waiting for the device\n", batch,
                        std::chrono::duration<double, std::micro>(t_enq - t_batch).count(), std::chrono::duration<double, std::micro>(t_done - t_enq).count());
            }
            if (hS->bb == 0.0) { tr_dump(); return zero_rhs(); }   // only reachable on the blind path (checked up front otherwise)
            // a NaN / Inf residual never satisfies rr <= tol^2 bb: stop here instead of running cg_max_iters garbage iterations
            // (the reference asserts in checkNanAndInf, src/systemHelpers.cpp:212-241)
            if (!std::isfinite(hS->rr) || !std::isfinite(hS->bb) || !std::isfinite(hS->rz)) {
                tr_dump();
                stats.cg_iters = hS->iters; stats.cg_converged = 0; stats.cg_rel_residual = NAN;
                err = "pressure solve: non-finite residual (NaN/Inf in the velocity field or right-hand side)";
                return 5;
            }
        }
        tr_dump();
        prev_iters = hS->iters;
        stats.cg_iters = hS->iters;
        stats.cg_rel_residual = std::sqrt(hS->rr / hS->bb);
        stats.cg_converged = hS->done ? 1 : 0;   // 0: stopped at cg_max_iters (callers that fix the iteration count expect that)
        return 0;
    }

    int grad_subtract(Real dt, int ext = 0) {
        const Real s = dt / (C.rho * C.h);
        { const Grid q = slab(ext);
          if (vec4()) LAUNCH((k_grad_subtract_v4<Real>), vec_blocks(q), 256, q, s, pres, U(0), U(1), U(2));
          else LAUNCH((k_grad_subtract<Real>), cell_blocks(q), 256, q, s, pres, U(0), U(1), U(2)); }
        CK(cudaGetLastError());
        maxv_valid = false;
        return 0;
    }
    // ext = rows of redundancy the caller needs on the projected velocity (viscosity must have produced ext + 0 rows;
    // the pressure is exchanged ext + 1 rows down / ext rows up after the solve)
    int pressure_impl(Real dt, int ext = 0) {
        const Real scale = (C.rho * C.h) / dt;
        if (vec4()) LAUNCH((k_divergence_v4<Real>), vec_blocks(g), 256, g, 2, scale, C.patm, U(0), U(1), U(2), rhs);
        else LAUNCH((k_divergence<Real>), cell_blocks(), 256, g, 2, scale, C.patm, U(0), U(1), U(2), rhs);
        CK(cudaGetLastError());
        if (!(P.cg_warm_start && have_q)) CK(cudaMemsetAsync(q, 0, (size_t)g.n * sizeof(Real), stream));
        if (int rc = cg_solve(rhs, q)) return rc;
        have_q = true;
        if (int rc = halo_exchange(q, ext + 1, ext)) return rc;
        LAUNCH((k_add_scalar<Real>), nblk(g.n), 256, g.n, C.patm, q, pres);
        CK(cudaGetLastError());
        return grad_subtract(dt, ext);
    }
    int pressure(double dt) override {
        if (int rc = pressure_impl((Real)(float)dt)) return rc;
        return allgather_velocity();
    }
    int pressure_given_p(double dt) override {
        if (int rc = grad_subtract((Real)(float)dt)) return rc;
        return allgather_velocity();
    }

    int poisson_solve(const void* b, void* p, int hd) override {
        if (int rc = check_hd(hd, err)) return rc;
        if (int rc = grid_in_any(b, rhs, hd)) return rc;
        LAUNCH((k_scale<Real>), nblk(g.n), 256, g.n, Real(-1), rhs);
        CK(cudaMemsetAsync(cg_d[0], 0, (size_t)g.n * sizeof(Real), stream));
        Real* x = curl[0];   // scratch: curl is rewritten by the next vorticity phase before use
        CK(cudaMemsetAsync(x, 0, (size_t)g.n * sizeof(Real), stream));
        if (int rc = cg_solve(rhs, x)) return rc;
        if (int rc = allgather_rows(x, sizeof(Real))) return rc;
        return grid_out_any(x, p, hd);
    }

    int vorticity_impl(Real dt) {
        { const Grid q = slab(1);
          if (vec4()) LAUNCH((k_curl_v4<Real>), vec_blocks(q), 256, q, U(0), U(1), U(2), curl[0], curl[1], curl[2], curl_n);
          else LAUNCH((k_curl<Real>), cell_blocks(q), 256, q, U(0), U(1), U(2), curl[0], curl[1], curl[2]); }
        CK(cudaMemsetAsync(d_scal, 0, sizeof(unsigned int), stream));
        if (vec4()) LAUNCH((k_vorticity_v4<Real>), vec_blocks(g), 256, g, dt, C.kvort, curl[0], curl[1], curl[2], curl_n, U(0), U(1), U(2), d_scal);
        else LAUNCH((k_vorticity<Real>), cell_blocks(), 256, g, dt, C.kvort, curl[0], curl[1], curl[2], U(0), U(1), U(2), d_scal);
        CK(cudaGetLastError());
        if (int rc = allreduce(d_scal, 1, ncclUint32, ncclMax)) return rc;
        maxv_valid = true;   // fused K0
        return 0;
    }
    // stand-alone hook: the velocity is valid on every row (previous all-gather), so curl's one-row redundancy is exact
    int vorticity(double dt) override {
        if (int rc = vorticity_impl((Real)(float)dt)) return rc;
        return allgather_velocity();
    }

    // System::updateVelocityField (src/waterSolver.cpp:41-96).  Multi-GPU: each phase runs on this rank's rows widened by the
    // rows the LATER stencil phases will read (forces 3, viscosity 2, projection 2, curl 1, confinement 0), then the
    // owners' rows are all-gathered once: every rank needs the whole field for its particle shard anyway.
    int water_phases(Real dt, bool timed) {
        if (int rc = advect_impl(dt, true, 3)) return rc;
        if (timed) CK(cudaEventRecord(ev[2], stream));
        if (int rc = forces_impl(dt, 3)) return rc;
        if (timed) CK(cudaEventRecord(ev[3], stream));
        if (int rc = viscosity_impl(dt, 2)) return rc;
        if (timed) CK(cudaEventRecord(ev[4], stream));
        if (int rc = pressure_impl(dt, 2)) return rc;
        if (timed) CK(cudaEventRecord(ev[5], stream));
        if (int rc = vorticity_impl(dt)) return rc;
        return allgather_velocity();
    }

    int velocity_field_update(double dt_) override { return water_phases((Real)(float)dt_, false); }

    int particles_impl(Real dt) {
        if (np == 0) return 0;
        CK(cudaMemsetAsync(seed, 0, g.n, stream));
        const bool fast = P.arith == INK_ARITH_FAST, uh = C.unit_h && g.n < (1ll << 31);
#define INK_K8(F, H) LAUNCH((k_particles_rk2<Real, F, H>), nblk(np), 256, g, C, dt, np, part[pcur], U(0), U(1), U(2), seed)
        if (tile_chunk > 0 && uh) {   // TMA-staged shared-memory tiles; upper bound on the work items, surplus blocks exit at once
            const unsigned int nb = (unsigned int)((np + tile_chunk - 1) / tile_chunk + 2ll * g.Y);
            if (fast) launch_k((k_particles_rk2_tile<Real, true>), nb, kTileThreads, (size_t)kTileSmemBytes, stream, g, C, dt, part[pcur], U(0), U(1), U(2), seed, tile_work, tile_nwork);
            else launch_k((k_particles_rk2_tile<Real, false>), nb, kTileThreads, (size_t)kTileSmemBytes, stream, g, C, dt, part[pcur], U(0), U(1), U(2), seed, tile_work, tile_nwork);
            ++stats.kernel_launches;
        } else if (fast) { if (uh) INK_K8(true, true); else INK_K8(true, false); }
        else { if (uh) INK_K8(false, true); else INK_K8(false, false); }
#undef INK_K8
        CK(cudaGetLastError());
        // K8 seeded the mask from the new positions, and the wall clamp put every particle inside the grid; phantom
        // cells created earlier persist in the reference's map (their velocity keeps entering calcTimeStep)
        seed_valid = true;
        seed_global = false;
        oob_now = false;
        ++since_sort;
        return 0;
    }
    int particles(double dt) override {
        if (int rc = particles_impl((Real)(float)dt)) return rc;
        if (P.resort_interval > 0 && since_sort >= P.resort_interval) return resort();
        return 0;
    }

    // System::solve (src/inkSolver.cpp:5-9)
    int step(double ttnr, double* dt_out) override {
        double dt_;
        CK(cudaEventRecord(ev[0], stream));
        if (int rc = calc_dt(ttnr, &dt_)) return rc;
        const Real dt = (Real)(float)dt_;   // `float timeStep` parameters narrow the double (Q15)
        CK(cudaEventRecord(ev[1], stream));
        if (int rc = water_phases(dt, true)) return rc;
        CK(cudaEventRecord(ev[6], stream));
        if (int rc = particles_impl(dt)) return rc;
        CK(cudaEventRecord(ev[7], stream));
        if (P.resort_interval > 0 && since_sort >= P.resort_interval)
            if (int rc = resort()) return rc;
        CK(cudaEventRecord(ev[8], stream));
        ev_valid = true;
        ++stats.steps;
        if (dt_out) *dt_out = dt_;
        return 0;
    }
    // System::checkNanAndInf (src/systemHelpers.cpp:212-241) on the device: out[0] = grid cells / particles with a NaN or Inf
    // component, out[1] = those whose velocity norm is not < 10000.  Multi-GPU: own rows and own particles, summed over ranks.
    int check_finite(int64_t out[2]) override {
        unsigned long long* d_chk = reinterpret_cast<unsigned long long*>(d_scal + 4);
        unsigned long long* h_chk = reinterpret_cast<unsigned long long*>(h_scal + 4);
        CK(cudaMemsetAsync(d_chk, 0, 2 * sizeof(unsigned long long), stream));
        const long long cells = (long long)(g.j1 - g.j0) * g.sy, base = (long long)g.j0 * g.sy;
        auto blocks = [](long long n) { return (unsigned int)std::min<long long>(kSMs * 8, std::max<long long>(1, (n + 255) / 256)); };
        LAUNCH((k_check_finite<Real>), blocks(cells), 256, cells, U(0) + base, U(1) + base, U(2) + base, 1, d_chk);   // old == curr between steps
        if (np > 0) {
            LAUNCH((k_check_finite<Real>), blocks(np), 256, np, part[pcur].p[0], part[pcur].p[1], part[pcur].p[2], 0, d_chk);
            LAUNCH((k_check_finite<Real>), blocks(np), 256, np, part[pcur].v[0], part[pcur].v[1], part[pcur].v[2], 1, d_chk);
        }
        CK(cudaGetLastError());
        if (int rc = allreduce(d_chk, 2, ncclUint64, ncclSum)) return rc;
        read_small(h_chk, d_chk, 2 * sizeof(unsigned long long));
        CK(cudaStreamSynchronize(stream));
        out[0] = (int64_t)h_chk[0]; out[1] = (int64_t)h_chk[1];
        return 0;
    }
    // phantom-cell state (Q8): part of what a checkpoint must carry for calcTimeStep to resume bit for bit
    int get_phantom(int32_t* have, double v[3]) override {
        *have = have_phantom ? 1 : 0;
        for (int a = 0; a < 3; ++a) v[a] = (double)phantom_v[a];
        return 0;
    }
    int set_phantom(int32_t have, const double v[3]) override {
        have_phantom = have != 0;
        for (int a = 0; a < 3; ++a) phantom_v[a] = (Real)v[a];
        return 0;
    }
    int phase_ms(float out[8]) override {
        if (!ev_valid) { err = "no ink_step yet"; return 2; }
        CK(cudaEventSynchronize(ev[8]));
        for (int a = 0; a < 8; ++a) CK(cudaEventElapsedTime(&out[a], ev[a], ev[a + 1]));
        return 0;
    }
};

// =================================================================================================
// ink_check_finite_noalloc — the checkNanAndInf reduction for callers that cannot allocate device memory.
//
// Every other entry point needs an ink_ctx, and ink_create cudaMallocs the whole simulation state.  Two kinds of host cannot
// go that way: (1) an application that reserved all of HBM for its own arena at start-up (a renderer, or a solver that already
// holds the grid and the particles in its own buffers): cudaMalloc simply fails there; (2) a frame loop that must not stall:
// cudaMalloc / cudaFree synchronise the whole device, so one call in the loop holds up every other stream of the process.
// Such a host keeps the state in ITS buffers and only wants this library to look at them.  The kernel below is
// k_check_finite (grid_kernels.cuh) with two changes: element t is read at x[t * stride] so that the AoS particle layout of
// the ABI (N x 3) works as well as three dense arrays, and the two counters are not atomically added to a cudaMalloc'ed
// accumulator: every block leaves its two partial counts in its own slot of the static array g_noalloc_scratch, and the block
// that takes the last ticket adds the slots up in slot order and writes the caller's output.  That array is all the scratch
// there is; its size bounds the number of blocks and with it the number of elements one call can cover (kNoAllocMaxN).
//
// The array lives in the module image: every GPU on which this library's kernels are loaded carries a copy of it until the
// process exits.  On the HOST the name g_noalloc_scratch is only a handle for the runtime — handing `g_noalloc_scratch`
// itself to a kernel would give it a host address — so the first call on each device (with that device current) asks the
// runtime for that device's copy with cudaGetSymbolAddress and keeps the answer in g_noalloc_dev[device].
// =================================================================================================
constexpr int kNoAllocBlock = 256;                       // threads per block
constexpr int kNoAllocPerThread = 32;                    // elements per thread: a block covers 8192 consecutive elements
constexpr int kNoAllocMaxBlocks = 32768;                 // slots in the scratch array
constexpr long long kNoAllocMaxN = (long long)kNoAllocMaxBlocks * kNoAllocBlock * kNoAllocPerThread;   // 2^28 elements per call
constexpr size_t kNoAllocHeadBytes = 256;                // [0, 256): the block ticket (must be 0 when a launch starts)
constexpr size_t kNoAllocScratchBytes = kNoAllocHeadBytes + (size_t)kNoAllocMaxBlocks * sizeof(uint2);   // 256 B + 256 KiB
static_assert(kNoAllocMaxN == INK_NOALLOC_MAX_N, "include/ink_capi.h states the limit");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

constexpr int kNoAllocMaxDevices = 64;
static void* g_noalloc_dev[kNoAllocMaxDevices] = {};     // device address of g_noalloc_scratch on device d (0 = not asked yet)
static std::mutex g_noalloc_mu;

template <class Real>
__global__ void __launch_bounds__(kNoAllocBlock) k_check_finite_noalloc(long long n, long long stride, const Real* __restrict__ x, const Real* __restrict__ y,
                                                                         const Real* __restrict__ z, int check_norm, unsigned char* __restrict__ scratch,
                                                                         unsigned long long* __restrict__ counts) {
    unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch);
    uint2* slot = reinterpret_cast<uint2*>(scratch + kNoAllocHeadBytes);
    __shared__ unsigned int s_bad[kNoAllocBlock / 32], s_big[kNoAllocBlock / 32];
    __shared__ bool s_last;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned int bad = 0, big = 0;
    const long long t0 = (long long)blockIdx.x * (kNoAllocBlock * kNoAllocPerThread) + threadIdx.x;
#pragma unroll 4
    for (int q = 0; q < kNoAllocPerThread; ++q) {
        const long long t = t0 + (long long)q * kNoAllocBlock;
        if (t < n) {
            const Real a = x[t * stride], b = y[t * stride], c = z[t * stride];
            if (!(isfinite(a) && isfinite(b) && isfinite(c))) ++bad;
            if (check_norm && !(norm3(a, b, c) < Real(10000))) ++big;     // the same two tests as k_check_finite
        }
    }
    bad = __reduce_add_sync(0xffffffffu, bad);
    big = __reduce_add_sync(0xffffffffu, big);
    if (lane == 0) { s_bad[wid] = bad; s_big[wid] = big; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int tb = 0, tg = 0;
        for (int q = 0; q < kNoAllocBlock / 32; ++q) { tb += s_bad[q]; tg += s_big[q]; }
        slot[blockIdx.x] = make_uint2(tb, tg);            // a slot is always written before it is read: nothing of an earlier call survives
        __threadfence();
        s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    unsigned long long sb = 0, sg = 0;
    for (unsigned int q = threadIdx.x; q < gridDim.x; q += kNoAllocBlock) {
        const uint2 v = __ldcg(&slot[q]);
        sb += v.x; sg += v.y;
    }
    __shared__ unsigned long long s_tot[2][kNoAllocBlock / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { sb += __shfl_xor_sync(0xffffffffu, sb, o); sg += __shfl_xor_sync(0xffffffffu, sg, o); }
    if (lane == 0) { s_tot[0][wid] = sb; s_tot[1][wid] = sg; }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long ab = 0, ag = 0;
        for (int q = 0; q < kNoAllocBlock / 32; ++q) { ab += s_tot[0][q]; ag += s_tot[1][q]; }
        counts[0] = ab; counts[1] = ag;
        *ticket = 0;                                       // leave the ticket as the next launch needs it
    }
}

static int noalloc_fail(int rc, const char* what, cudaError_t e = cudaSuccess) {
    char buf[512];
    if (e != cudaSuccess) snprintf(buf, sizeof buf, "ink_check_finite_noalloc: %s -> %s", what, cudaGetErrorString(e));
    else snprintf(buf, sizeof buf, "ink_check_finite_noalloc: %s", what);
    g_create_error = buf;
    return rc;
}

static int check_finite_noalloc(const void* x, const void* y, const void* z, int64_t n, int64_t stride, int dtype, int check_norm,
                                uint64_t* counts_dev, void* cuda_stream) {
    if (!counts_dev) return noalloc_fail(2, "counts_dev is NULL");
    if (n < 0) return noalloc_fail(2, "n is negative");
    if (n > kNoAllocMaxN) {
        char buf[160];
        snprintf(buf, sizeof buf, "n = %lld is beyond INK_NOALLOC_MAX_N = %lld elements per call (the static scratch array has one slot per 8192 elements): split the range",
                 (long long)n, kNoAllocMaxN);
        return noalloc_fail(2, buf);
    }
    if (n > 0 && (!x || !y || !z)) return noalloc_fail(2, "x, y or z is NULL");
    if (stride < 1) return noalloc_fail(2, "stride must be >= 1 (1: three dense arrays, 3: one N x 3 array passed as p, p + 1, p + 2)");
    if (dtype != 32 && dtype != 64) return noalloc_fail(2, "dtype must be 32 (float) or 64 (double)");
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { cudaGetLastError(); return noalloc_fail(3, "cudaGetDevice (no CUDA device: there is no CPU path)", e); }
    if (dev < 0 || dev >= kNoAllocMaxDevices) return noalloc_fail(3, "device ordinal beyond 63");
    void* scratch;
    {
        // once per device, with that device current: where this device keeps its copy of g_noalloc_scratch
        std::lock_guard<std::mutex> lk(g_noalloc_mu);
        if (!g_noalloc_dev[dev]) {
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p) { cudaGetLastError(); return noalloc_fail(1, "cudaGetSymbolAddress(g_noalloc_scratch)", e); }
            g_noalloc_dev[dev] = p;
        }
        scratch = g_noalloc_dev[dev];
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    // the ticket has to be 0 when the kernel starts; a launch that ran to its end left it so, this covers one that did not
    e = cudaMemsetAsync(scratch, 0, kNoAllocHeadBytes, st);
    if (e != cudaSuccess) return noalloc_fail(1, "cudaMemsetAsync(ticket)", e);
    const unsigned int grid = (unsigned int)std::max<long long>(1, (n + kNoAllocBlock * kNoAllocPerThread - 1) / (kNoAllocBlock * kNoAllocPerThread));
    unsigned long long* out = reinterpret_cast<unsigned long long*>(counts_dev);
    if (dtype == 32)
        k_check_finite_noalloc<float><<<grid, kNoAllocBlock, 0, st>>>((long long)n, (long long)stride, (const float*)x, (const float*)y, (const float*)z, check_norm,
                                                                      (unsigned char*)scratch, out);
    else
        k_check_finite_noalloc<double><<<grid, kNoAllocBlock, 0, st>>>((long long)n, (long long)stride, (const double*)x, (const double*)y, (const double*)z, check_norm,
                                                                       (unsigned char*)scratch, out);
    e = cudaGetLastError();
    if (e != cudaSuccess) return noalloc_fail(1, "kernel launch", e);
    return 0;
}

}  // namespace ink

// -------------------------------------------------------------------------------------------------
using namespace ink;
struct ink_ctx { CtxBase* impl; };

extern "C" {

void ink_default_params(ink_params* p) {
    memset(p, 0, sizeof *p);
    p->X = 20; p->Y = 45; p->Z = 20;           // src/system.h:15-17
    p->cell_dim = 1.f; p->density = .95f; p->k_vort = 1.f; p->viscosity = 1.0016f; p->p_atm = 1.f;
    p->gravity[0] = 0.f; p->gravity[1] = -0.1f; p->gravity[2] = 0.f;
    p->k_cfl = 0.2f; p->dt_min = 0.005f; p->dt_max = 1.f; p->buffer_size = 3;
    p->precision = 32; p->arith = INK_ARITH_STRICT; p->device = 0;
    p->cg_rtol = 1e-6; p->cg_max_iters = 20000; p->cg_check_every = 16; p->cg_warm_start = 1; p->cg_precond = INK_PCG_MG_TRI;
    p->resort_interval = 16;
    p->rank = 0; p->world_size = 1;
}

int ink_nccl_unique_id(uint8_t out[128]) {
    memset(out, 0, 128);
    const std::string e = nccl().load();
    if (!e.empty()) { g_create_error = e; return 4; }
    ncclUniqueId id;
    if (nccl().GetUniqueId(&id) != ncclSuccess) { g_create_error = "ncclGetUniqueId failed"; return 4; }
    memcpy(out, &id, sizeof id);
    return 0;
}

int ink_create(const ink_params* p, ink_ctx** out) {
    if (!p || !out) return 2;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        g_create_error = "no CUDA device: libinkb200 has no CPU fallback";
        return 3;
    }
    CtxBase* c;
    if (p->precision == 32) c = new Ctx<float>();
    else if (p->precision == 64) c = new Ctx<double>();
    else { g_create_error = "precision must be 32 or 64"; return 2; }
    c->P = *p;
    if (c->P.precision == 32 && c->P.cg_rtol < 1e-7) c->P.cg_rtol = 1e-7;
    int rc = c->init();
    if (rc) { g_create_error = c->err; delete c; return rc; }
    *out = new ink_ctx{c};
    return 0;
}
void ink_destroy(ink_ctx* c) {
    if (!c) return;
    delete c->impl;
    delete c;
}
const char* ink_last_error(const ink_ctx* c) { return c ? c->impl->err.c_str() : g_create_error.c_str(); }
int ink_sync(ink_ctx* c) { return cudaStreamSynchronize(c->impl->stream) == cudaSuccess ? 0 : 1; }

int ink_set_particles(ink_ctx* c, int drop, const void* pos, const void* vel, int64_t n, int hd) { return c->impl->set_particles(drop, pos, vel, n, hd); }
int ink_set_particles_shard(ink_ctx* c, int drop, const void* pos, const void* vel, int64_t n, int hd) { return c->impl->set_particles_shard(drop, pos, vel, n, hd); }
int ink_num_drops(const ink_ctx* c) { return c->impl->num_drops(); }
int64_t ink_num_particles(const ink_ctx* c, int drop) { return c->impl->num_particles(drop); }
int ink_get_particles(ink_ctx* c, int drop, void* pos, void* vel, int hd) { return c->impl->get_particles(drop, pos, vel, hd); }
int ink_set_velocity(ink_ctx* c, const void* u, const void* v, const void* w, int hd) { return c->impl->set_velocity(u, v, w, hd); }
int ink_get_field(ink_ctx* c, int which, void* u, void* v, void* w, int hd) { return c->impl->get_field(which, u, v, w, hd); }
int ink_get_pressure(ink_ctx* c, void* p, int hd) { return c->impl->get_pressure(p, hd); }
int ink_set_pressure(ink_ctx* c, const void* p, int hd) { return c->impl->set_pressure(p, hd); }
int ink_get_warm_start(ink_ctx* c, void* q, int hd) { return c->impl->get_warm_start(q, hd); }
int ink_set_warm_start(ink_ctx* c, const void* q, int hd) { return c->impl->set_warm_start(q, hd); }
int ink_get_mark(ink_ctx* c, uint8_t* m) { return c->impl->get_mark(m); }
int ink_get_divergence(ink_ctx* c, void* d, int hd) { return c->impl->get_divergence(d, hd); }
int ink_sample_velocity(ink_ctx* c, const void* pos, int64_t n, void* out, int hd) { return c->impl->sample_velocity(pos, n, out, hd); }

int ink_step(ink_ctx* c, double ttnr, double* dt_out) { return c->impl->step(ttnr, dt_out); }
int ink_calc_dt(ink_ctx* c, double ttnr, double* dt_out) { return c->impl->calc_dt(ttnr, dt_out); }
int ink_advect(ink_ctx* c, double dt) { return c->impl->advect(dt); }
int ink_forces(ink_ctx* c, double dt) { return c->impl->forces(dt); }
int ink_viscosity(ink_ctx* c, double dt) { return c->impl->viscosity(dt); }
int ink_pressure_rhs(ink_ctx* c, double dt, void* b, int hd) { return c->impl->pressure_rhs(dt, b, hd); }
int ink_pressure(ink_ctx* c, double dt) { return c->impl->pressure(dt); }
int ink_pressure_given_p(ink_ctx* c, double dt) { return c->impl->pressure_given_p(dt); }
int ink_poisson_solve(ink_ctx* c, const void* b, void* p, int hd) { return c->impl->poisson_solve(b, p, hd); }
int ink_vorticity(ink_ctx* c, double dt) { return c->impl->vorticity(dt); }
int ink_velocity_field_update(ink_ctx* c, double dt) { return c->impl->velocity_field_update(dt); }
int ink_particles(ink_ctx* c, double dt) { return c->impl->particles(dt); }
int ink_resort_particles(ink_ctx* c) { return c->impl->resort(); }
int ink_prefetch_positions(ink_ctx* c, int drop, const void* pos, int64_t n, int hd) { return c->impl->prefetch_positions(drop, pos, n, hd); }
int ink_commit_positions(ink_ctx* c, int drop) { return c->impl->commit_positions(drop); }
int ink_fetch_positions_async(ink_ctx* c, int drop, void* pos, int hd) { return c->impl->fetch_positions_async(drop, pos, hd); }
int ink_sync_io(ink_ctx* c) { return c->impl->sync_io(); }

int ink_seed_emitter(ink_ctx* c, int drop, const float* radius, int nr, const float* sin_i, const float* cos_i, int ni, const float* sin_j, const float* cos_j, int nj) {
    return c->impl->seed_emitter(drop, radius, nr, sin_i, cos_i, ni, sin_j, cos_j, nj);
}
int ink_seed_box(ink_ctx* c, int drop, const int32_t* rand_xyz, int64_t n) { return c->impl->seed_box(drop, rand_xyz, n); }
int ink_check_finite(ink_ctx* c, int64_t out[2]) { return c->impl->check_finite(out); }
int ink_get_phantom(ink_ctx* c, int32_t* have, double v[3]) { return c->impl->get_phantom(have, v); }
int ink_set_phantom(ink_ctx* c, int32_t have, const double v[3]) { return c->impl->set_phantom(have, v); }
int ink_get_stream(ink_ctx* c, void** cuda_stream) { *cuda_stream = (void*)c->impl->stream; return 0; }
int ink_get_stats(ink_ctx* c, ink_stats* out) { *out = c->impl->stats; return 0; }
int ink_get_phase_ms(ink_ctx* c, float out[8]) { return c->impl->phase_ms(out); }
int ink_check_finite_noalloc(const void* x, const void* y, const void* z, int64_t n, int64_t stride, int dtype, int check_norm, uint64_t* counts_dev,
                             void* cuda_stream) {
    return check_finite_noalloc(x, y, z, n, stride, dtype, check_norm, counts_dev, cuda_stream);
}

}  // extern "C"
