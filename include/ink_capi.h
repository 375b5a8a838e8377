/*
 * ink_capi.h — C ABI of the B200-native Ink-Simulation hot path (libinkb200.so).
 *
 * The reference (zhang-tianran/Ink-Simulation) has no FFI; its seam is the C++ class `System`
 * (src/system.h:70-138).  This ABI sits *underneath* a drop-in `System` (host/system.h in this
 * repo): every entry point names the reference member it replaces.  Plain pointers and sizes only;
 * no C++/Eigen/torch types cross the boundary.  All calls return 0 on success, non-zero on error
 * (message via ink_last_error).  One ctx = one host thread = one CUDA device; all work is ordered
 * on the ctx's OWN non-blocking CUDA stream (ink_get_stream), never the legacy NULL stream.  Consequences for DEVICE
 * pointers: inputs must be complete before the call (the library's stream is not ordered against the caller's streams —
 * synchronise, or make the ctx stream wait on your event via ink_get_stream); device outputs and the caller's input
 * buffers are complete / free again when the call returns (the streaming calls ink_prefetch_positions /
 * ink_fetch_positions_async are the documented exception: they return at once).  There is NO CPU fallback:
 * ink_create fails if no sm_100 device is present.
 *
 * Array conventions (identical to oracle/ and the compiled-reference shim):
 *   grid fields  : dense (X, Y, Z) C-order, z fastest: idx = (i*Y + j)*Z + k
 *   particles    : AoS xyz, N x 3 (the layout of Particle::position / ::velocity, src/system.h:52-60)
 *   host_dtype   : 32 = float arrays, 64 = double arrays (independent of the ctx's compute precision)
 * Pointers may be host (pageable or pinned) or device pointers; the library detects which.
 */
#ifndef INK_CAPI_H
#define INK_CAPI_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ink_ctx ink_ctx;

enum { INK_ARITH_STRICT = 0, INK_ARITH_FAST = 1 };
enum { INK_FIELD_OLD = 0, INK_FIELD_CURR = 1, INK_FIELD_CURL = 2 };
/* INK_PCG_MG: V(2,2), piecewise-constant transfer (csrc/mg.cuh); INK_PCG_MG_TRI (default): V(1,1), trilinear transfer
 * (csrc/mg_tri.cuh) — fewer CG iterations and less than half the passes per iteration */
enum { INK_PCG_NONE = 0, INK_PCG_MG = 1, INK_PCG_MG_TRI = 2 };

/* Run-time copy of the reference's compile-time globals (src/system.h:13-38) + solver knobs. */
typedef struct ink_params {
    int32_t X, Y, Z;        /* WATERGRID_X/Y/Z            src/system.h:15-17 */
    float cell_dim;         /* CELL_DIM                   :22 */
    float density;          /* DENSITY                    :25 */
    float k_vort;           /* K_VORT                     :27 */
    float viscosity;        /* VISCOSITY                  :30 */
    float p_atm;            /* ATMOSPHERIC_PRESSURE       :31 */
    float gravity[3];       /* gravity                    :34 */
    float k_cfl;            /* K_CFL                      :36 */
    float dt_min;           /* MIN_TIMESTEP               :37 */
    float dt_max;           /* MAX_TIMESTEP               :38 */
    int32_t buffer_size;    /* BUFFER_SIZE                :23 */
    int32_t precision;      /* 32 (reference arithmetic) or 64 */
    int32_t arith;          /* INK_ARITH_STRICT: reference operation order, no FMA contraction (bit-identical to the
                               compiled reference in fp32); INK_ARITH_FAST: FMA lerp form in the particle kernel */
    int32_t device;         /* CUDA device ordinal */
    /* Pressure solve: replaces Eigen::SparseLU (src/system.h:105, src/waterSolver.cpp:259,279) by matrix-free CG */
    double  cg_rtol;        /* stop when ||r||_2 <= cg_rtol * ||b||_2 */
    int32_t cg_max_iters;
    int32_t cg_check_every; /* iterations between host convergence polls */
    int32_t cg_warm_start;  /* 1: start from the previous step's pressure */
    int32_t cg_precond;     /* INK_PCG_NONE | INK_PCG_MG | INK_PCG_MG_TRI */
    int32_t resort_interval;/* re-bin particles by cell every this many particle updates (0 = never) */
    /* multi-GPU (one process per GPU).  world_size==1: single device. */
    int32_t rank, world_size;
    uint8_t nccl_id[128];   /* ncclUniqueId from ink_nccl_unique_id on rank 0, broadcast by the caller */
} ink_params;

typedef struct ink_stats {
    double  dt;             /* last time step */
    int32_t cg_iters;       /* iterations of the last pressure solve */
    double  cg_rel_residual;/* ||r||/||b|| at exit (recurrence residual) */
    int32_t advect_skipped; /* 1 if the convection phase was proven to be the identity and skipped */
    int64_t steps;          /* ink_step calls so far */
    int64_t resorts;        /* particle re-binnings so far */
    int64_t kernel_launches;/* kernels executed for this ctx so far (kernel nodes of replayed CUDA graphs included) */
    int32_t cg_converged;   /* 1: the last pressure solve met cg_rtol; 0: it stopped at cg_max_iters (no error: callers that fix the
                               iteration count rely on it).  A NON-FINITE residual is an error: ink_step / ink_pressure return non-zero */
    int32_t graph_launches; /* CUDA-graph launches by this ctx so far (each replays one captured PCG iteration) */
} ink_stats;

void ink_default_params(ink_params* p);                 /* the committed values of src/system.h:13-38 */
int  ink_nccl_unique_id(uint8_t out[128]);
int  ink_create(const ink_params* p, ink_ctx** out);    /* System::System + initWaterGrid (src/system.cpp:8,40-60); no LU setup */
void ink_destroy(ink_ctx* c);
const char* ink_last_error(const ink_ctx* c);           /* c may be NULL: last create error */
int  ink_sync(ink_ctx* c);

/* ---- state in / out ------------------------------------------------------------------------ */
/* m_ink[drop] = particles (src/system.h:115; initFromFile-style seeding src/system.cpp:117-148).  vel may be NULL (zero).
 * Multi-GPU: every rank passes the FULL drop; each rank keeps its contiguous index shard. */
int  ink_set_particles(ink_ctx* c, int drop, const void* pos_xyz, const void* vel_xyz, int64_t n, int host_dtype);
/* Same, but the caller hands over only THIS rank's particles of the drop (a per-rank loader); identical to
 * ink_set_particles on one GPU. */
int  ink_set_particles_shard(ink_ctx* c, int drop, const void* pos_xyz, const void* vel_xyz, int64_t n_local, int host_dtype);
/* Particle seeding ON THE DEVICE.  What is sequential or libm-dependent in the reference's seeders stays with the
 * caller as small tables; every per-particle operation runs in a kernel, in the reference's expression order, and the particle
 * arrays never cross PCIe.  Results equal the compiled reference's arrays bit for bit (tests/test_host_cpp.py, -m gpu).
 * ink_seed_emitter: the 120-radius emitParticleHemisphere scene of System::init (src/system.cpp:18-23, src/inkSolver.cpp:11-42);
 *   radius[nr] / the azimuth loop values' sinf, cosf [ni] / the polar loop values' sinf, cosf [nj] as the reference's FLOAT loop
 *   counters produce them; particle (r*ni + i)*nj + j.  ink_seed_box: initParticles (src/system.cpp:62-64,93-115) from the raw
 *   rand() stream, 3 draws (x, y, z) per particle.  Multi-GPU: every rank passes the full tables / stream, keeps its index shard. */
int  ink_seed_emitter(ink_ctx* c, int drop, const float* radius, int nr, const float* sin_i, const float* cos_i, int ni,
                      const float* sin_j, const float* cos_j, int nj);
int  ink_seed_box(ink_ctx* c, int drop, const int32_t* rand_xyz, int64_t n);
int  ink_num_drops(const ink_ctx* c);
int64_t ink_num_particles(const ink_ctx* c, int drop);  /* local shard size under multi-GPU */
/* getInkParticles() (src/system.cpp:151-153): original particle order; either pointer may be NULL. */
int  ink_get_particles(ink_ctx* c, int drop, void* pos_xyz, void* vel_xyz, int host_dtype);
/* oldVelocity = currVelocity = (u,v,w)  (Cell, src/system.h:41-50) */
int  ink_set_velocity(ink_ctx* c, const void* u, const void* v, const void* w, int host_dtype);
/* getWaterGrid() (src/system.cpp:155-157) as three dense arrays; which = INK_FIELD_* */
int  ink_get_field(ink_ctx* c, int which, void* u, void* v, void* w, int host_dtype);
int  ink_get_pressure(ink_ctx* c, void* p, int host_dtype);
int  ink_set_pressure(ink_ctx* c, const void* p, int host_dtype);   /* warm-start seed / pressure_given tests */
/* The Poisson solver's own unknown q = p - p_atm exactly as it is held (the start vector of the next solve when
 * cg_warm_start = 1).  For checkpoints: p = q + p_atm has lost q's low bits, so a run restored through ink_set_pressure
 * continues within the solver tolerance, one restored through ink_set_warm_start continues bit for bit.  No reference
 * equivalent (SparseLU has no state); SURVEY 8f-3 "matching load". */
int  ink_get_warm_start(ink_ctx* c, void* q, int host_dtype);
int  ink_set_warm_start(ink_ctx* c, const void* q, int host_dtype);
int  ink_get_mark(ink_ctx* c, uint8_t* mark);           /* force mask of the last forces phase */
int  ink_get_divergence(ink_ctx* c, void* div, int host_dtype);  /* getDivergence on oldVelocity, src/systemHelpers.cpp:30-43 */
int  ink_sample_velocity(ink_ctx* c, const void* pos_xyz, int64_t n, void* out_xyz, int host_dtype); /* getVelocity :110-115 */

/* ---- the hot path --------------------------------------------------------------------------- */
/* System::solve (src/inkSolver.cpp:5-9): updateWaterGrid then updateParticles; *dt_out = returned time step. */
int  ink_step(ink_ctx* c, double time_to_next_render, double* dt_out);
/* Per-phase hooks, each = the phase followed by the reference's `old = curr` sweep (src/waterSolver.cpp:41-96). */
int  ink_calc_dt(ink_ctx* c, double time_to_next_render, double* dt_out);  /* calcTimeStep      src/waterSolver.cpp:13-38 */
int  ink_advect(ink_ctx* c, double dt);                 /* applyConvection    :109-128 */
int  ink_forces(ink_ctx* c, double dt);                 /* applyExternalForces:158-191 */
int  ink_viscosity(ink_ctx* c, double dt);              /* applyViscosity     :225-238 */
int  ink_pressure_rhs(ink_ctx* c, double dt, void* b_out, int host_dtype); /* RHS of calculatePressure :264-278 (no state change) */
int  ink_pressure(ink_ctx* c, double dt);               /* applyPressure      :284-296 (RHS + solve + gradient subtract) */
int  ink_pressure_given_p(ink_ctx* c, double dt);       /* gradient subtract with the pressure set by ink_set_pressure */
int  ink_poisson_solve(ink_ctx* c, const void* b, void* p, int host_dtype); /* llt.solve(b) :279 for arbitrary b */
int  ink_vorticity(ink_ctx* c, double dt);              /* applyVorticity     :203-222 */
int  ink_velocity_field_update(ink_ctx* c, double dt);  /* updateVelocityField:41-96 */
int  ink_particles(ink_ctx* c, double dt);              /* updateParticles    src/inkSolver.cpp:44-84 */
int  ink_resort_particles(ink_ctx* c);                  /* re-bin now (no reference equivalent; layout only) */

/* ---- streaming particle I/O (no reference equivalent: the reference keeps m_ink in host memory) ------------------
 * For callers that feed positions from the host every step and read them back every step (the per-frame .ply writer of
 * src/inkSim.cpp:45-96, bench.py's e2e leg).  Copies run on two dedicated copy streams with double-buffered device
 * staging, so the H2D of the next input and the D2H of the last result overlap the ink_step in between:
 *     ink_prefetch_positions(next)  ->  ink_commit_positions  ->  ink_step  ->  ink_fetch_positions_async(out)
 * Buffers should be pinned host memory (pageable memory works but serialises).  ink_commit_positions replaces the
 * positions of an existing drop (same count; binning and order kept), i.e. ink_set_particles without re-binning. */
int  ink_prefetch_positions(ink_ctx* c, int drop, const void* pos_xyz, int64_t n, int host_dtype);
int  ink_commit_positions(ink_ctx* c, int drop);
int  ink_fetch_positions_async(ink_ctx* c, int drop, void* pos_xyz, int host_dtype);
int  ink_sync_io(ink_ctx* c);   /* all outstanding copies and the compute stream are done */

/* System::checkNanAndInf (src/systemHelpers.cpp:212-241) as one device reduction: out[0] = grid cells + particles with a NaN / Inf
 * component (velocity of every cell; position and velocity of every particle), out[1] = cells + particles whose velocity norm
 * is not < 10000.  Both 0 = the reference's asserts would pass.  8 bytes cross PCIe, not the state.  Multi-GPU: collective. */
int  ink_check_finite(ink_ctx* c, int64_t out[2]);
/* Phantom-cell state (out-of-grid particles create map cells that only ever receive dt*g and enter calcTimeStep's
 * maximum, src/waterSolver.cpp:13-38,158-191): have = any such cell exists, v = their common velocity.  For checkpoints. */
int  ink_get_phantom(ink_ctx* c, int32_t* have, double v[3]);
int  ink_set_phantom(ink_ctx* c, int32_t have, const double v[3]);
/* The ctx's cudaStream_t (as void*): record events on it to time the library, or make it wait on yours.  No reference equivalent. */
int  ink_get_stream(ink_ctx* c, void** cuda_stream);
int  ink_get_stats(ink_ctx* c, ink_stats* out);
/* Device-time of every phase of the last ink_step in ms: dt, advect, forces, viscosity, pressure, vorticity, particles, resort */
int  ink_get_phase_ms(ink_ctx* c, float out[8]);

/* ---- for callers that cannot allocate device memory --------------------------------------------------------------
 * This entry point has no counterpart in the reference (which has no device at all); it is ink_check_finite's reduction
 * without an ink_ctx.  Every other call needs a context, and ink_create cudaMallocs the whole simulation state.  That
 * rules out a host application which handed all of HBM to its own arena at start-up (cudaMalloc fails there), and a
 * frame loop that may not stall (cudaMalloc and cudaFree synchronise the device and hold up every other stream of the
 * process).  Such a host keeps grid and particles in its own buffers; this call inspects them in place and allocates
 * nothing: no cudaMalloc / cudaFree, no pinned memory, no library that allocates behind the scenes.
 *
 *   x, y, z      DEVICE pointers (memory of the CURRENT device): element t is (x[t*stride], y[t*stride], z[t*stride]).
 *                stride = 1: three dense arrays (a velocity field).  stride = 3 with (p, p + 1, p + 2): one N x 3 array,
 *                the layout of Particle::position / ::velocity used by ink_set_particles.
 *   n            number of elements, 0 <= n <= INK_NOALLOC_MAX_N; anything larger is refused (return 2): split the range.
 *   dtype        32 = float, 64 = double.
 *   check_norm   1: also count elements whose norm() is not < 10000 (velocities); 0: positions (NaN / Inf test only).
 *   counts_dev   DEVICE pointer to two uint64, OVERWRITTEN (not added to): [0] = elements with a NaN / Inf component,
 *                [1] = elements failing the norm test (0 when check_norm = 0).  Same counting rules as ink_check_finite.
 *   cuda_stream  the cudaStream_t (of the current device) to run on, NULL = the legacy default stream.  The call only
 *                enqueues: the counts are there when work queued on that stream before them has finished.
 *
 * The kernel's only scratch memory is a file-scope device array of 256 B + 256 KiB inside the library image (one slot
 * of partial counts per block of 8192 elements, hence the limit).  Every GPU on which the library's kernels get
 * loaded holds one copy of it until the process exits.  What follows from that:
 *   - the FIRST call on each device looks up that device's copy (cudaGetSymbolAddress) with the caller's current device;
 *     make that first call during set-up, and in any case before a stream capture is begun on the stream;
 *   - all calls on one device share the array.  Two calls whose kernels can run at the same time on the GPU (different
 *     streams, or different host threads on different streams) race on it and give wrong counts; calls on ONE stream,
 *     or otherwise ordered one after the other, are safe.  Nothing is carried from call to call: the block ticket is
 *     cleared on the stream before each launch, and every slot is written before it is read.
 * Returns 0, or non-zero with a message from ink_last_error(NULL) on the calling thread. */
#define INK_NOALLOC_MAX_N 268435456LL   /* 2^28 elements per call */
int  ink_check_finite_noalloc(const void* x, const void* y, const void* z, int64_t n, int64_t stride, int dtype, int check_norm,
                              uint64_t* counts_dev, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* INK_CAPI_H */
