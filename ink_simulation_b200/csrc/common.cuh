// common.cuh — shared device/host helpers for the B200 (sm_100a) Ink-Simulation hot path.
//
// Internal grid layout: SoA components, y SLOWEST, z FASTEST:  idx(i,j,k) = (j*X + i)*Z + k.
// (The public C ABI speaks the reference-shim order (i*Y + j)*Z + k; ink_capi converts on device.)
// y-slowest makes a y-slab a contiguous range and a halo plane X*Z contiguous values (multi-GPU).
//
// The whole translation unit is compiled with -fmad=false: plain C++ expressions below therefore
// round exactly like the reference's x86-64 (no FMA) build, operation for operation.  Kernels that
// want FMAs (INK_ARITH_FAST particle path, CG) call fma()/fmaf() explicitly.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// host+device helpers: the multigrid level operators are also run on the HOST by tests/mg_host_harness.cu
#define INK_HD __host__ __device__ __forceinline__

namespace ink {

struct Grid {
    int X, Y, Z;
    int sx;          // stride of i  (= Z)
    long long sy;    // stride of j  (= X*Z)
    long long n;     // X*Y*Z
    int j0, j1;      // y-range this rank updates (whole grid on one GPU)
};

// Reference globals (src/system.h:13-38) in the compute precision, pre-converted on the host the way
// the reference's expressions convert them (float expressions first, then widened).
template <class Real>
struct Consts {
    Real h;          // CELL_DIM
    Real bx, by, bz; // (Real)(float)(X*CELL_DIM) ... upper bounds of isInBounds (src/systemHelpers.cpp:141-147)
    Real rho;        // DENSITY
    Real kvort;      // K_VORT
    Real visc;       // VISCOSITY
    Real patm;       // ATMOSPHERIC_PRESSURE
    Real g[3];       // gravity
    Real lo;         // (Real)0.01f
    Real cx, cy, cz; // (Real)(float)(X*CELL_DIM - 0.01f)
    int unit_h;      // CELL_DIM == 1: x / h is the identity, skip the IEEE division
    int R;           // BUFFER_SIZE
};

__device__ __forceinline__ long long cidx(const Grid& g, int i, int j, int k) {
    return (long long)j * g.sy + (long long)i * g.sx + k;
}
__device__ __forceinline__ bool in_idx(const Grid& g, int i, int j, int k) {  // isInBoundsbyIdx, src/systemHelpers.cpp:150-155
    return (i >= 0) && (i < g.X) && (j >= 0) && (j < g.Y) && (k >= 0) && (k < g.Z);
}

// Eigen 3.4 Vector3f::norm(): sqrt(x*x + (y*y + z*z)) (redux unroller splits 3 coefficients 1 + 2).
__device__ __forceinline__ float norm3(float x, float y, float z) { return sqrtf(x * x + (y * y + z * z)); }
__device__ __forceinline__ double norm3(double x, double y, double z) { return sqrt(x * x + (y * y + z * z)); }

__device__ __forceinline__ float rfloor(float x) { return floorf(x); }
__device__ __forceinline__ double rfloor(double x) { return floor(x); }
__device__ __forceinline__ float rfma(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double rfma(double a, double b, double c) { return fma(a, b, c); }

// streaming (read-once / write-once) accesses: keep them out of L1 so the gathered grid stays there
template <class T>
__device__ __forceinline__ T ld_stream(const T* p) { return __ldcs(p); }
template <class T>
__device__ __forceinline__ void st_stream(T* p, T v) { __stcs(p, v); }

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

constexpr int kSMs = 148;  // B200

// spin until a sequence number published by a PEER GPU (protocol: cg.cuh) has reached `want`; then order the payload reads after it
__device__ __forceinline__ void p2p_spin(const volatile unsigned long long* flag, unsigned long long want) {
    while (*flag < want) {
#if defined(__CUDA_ARCH__) && __CUDA_ARCH__ >= 700
        __nanosleep(64);
#endif
    }
    __threadfence_system();
}

// Programmatic dependent launch (sm_90+).  The step is a chain of ~300 short kernels on one stream; with plain launches
// each pays the full launch + drain latency (2-3 us, which is most of a coarse multigrid level's kernel).  Every kernel
// of the library therefore starts with INK_PDL(): griddepcontrol.wait blocks until the PRECEDING kernel has completed
// and its memory is visible (a no-op for a normally launched kernel), then launch_dependents lets the NEXT kernel's
// blocks be scheduled while this one runs (they sit in their own wait).  Nothing touches global memory before the
// wait, so RAW and WAR hazards between consecutive kernels are ordered exactly as with plain stream serialization.
__device__ __forceinline__ void pdl_enter() {
#if defined(__CUDA_ARCH__) && __CUDA_ARCH__ >= 900
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}
#define INK_PDL() ::ink::pdl_enter()

}  // namespace ink
