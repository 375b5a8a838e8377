// mg_host_harness.cu — TEST INFRASTRUCTURE.  Runs the multigrid level operators of csrc/mg_tri.cuh on the HOST (they are
// __host__ __device__ functions; nvcc builds this file as a plain CPU program) so tests/test_mg_host.py can check the
// V-cycle against the numpy prototype (scripts/mg_proto.py) in a container without a GPU.
//   usage: mg_host_harness X Y Z j0 j1 prec(32|64) threads in.bin out.bin
// in.bin: X*Y*Z reals of `prec` in the INTERNAL layout (y slowest, z fastest); out.bin: z = B r on rows [j0, j1), 0 elsewhere.
// `threads` emulates the (tid, nth) striding of a launch: every operator is run for tid = 0..threads-1.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../ink_simulation_b200/csrc/mg_tri.cuh"

using namespace ink;

template <class Real>
static int run(MgDims d0, int T, const char* in, const char* out) {
    std::vector<MgDims> d{d0};
    while ((long long)d.back().X * (d.back().j1 - d.back().j0) * d.back().Z > kMg3CoarsestCells) {
        const MgDims& q = d.back();
        d.push_back(MgDims{(q.X + 1) / 2, (q.Y + 1) / 2, (q.Z + 1) / 2, q.j0 >> 1, (q.j1 + 1) >> 1});
    }
    const int L = (int)d.size();
    std::vector<std::vector<Real>> r(L), x(L), t(L);
    for (int l = 0; l < L; ++l) {
        const size_t n = (size_t)d[l].X * d[l].Y * d[l].Z;
        r[l].assign(n, Real(0)); x[l].assign(n, Real(0)); t[l].assign(n, Real(0));
    }
    FILE* f = fopen(in, "rb");
    if (!f || fread(r[0].data(), sizeof(Real), r[0].size(), f) != r[0].size()) { fprintf(stderr, "cannot read %s\n", in); return 2; }
    fclose(f);
    Real s = 1;
    const bool fused = getenv("MG_HARNESS_FUSED") && atoi(getenv("MG_HARNESS_FUSED")) != 0;   // the one-pass down operator (levels >= 1 on the GPU)
    for (int l = 0; l + 1 < L; ++l) {
        if (fused) {
            for (int tid = 0; tid < T; ++tid) mg3_down_fused<Real>(d[l], d[l + 1], r[l].data(), r[l + 1].data(), tid, T);
        } else {
            for (int tid = 0; tid < T; ++tid) mg3_residual<Real>(d[l], r[l].data(), t[l].data(), tid, T);
            for (int tid = 0; tid < T; ++tid) mg3_restrict<Real>(d[l], d[l + 1], t[l].data(), r[l + 1].data(), tid, T);
        }
        s *= 2;
    }
    {
        const int q = L - 1;
        const Real cj = Real(kMgOmega) / (Real(6) * s);
        for (int j = d[q].j0; j < d[q].j1; ++j)
            for (int i = 0; i < d[q].X; ++i)
                for (int k = 0; k < d[q].Z; ++k) { const long long c = mg_idx(d[q], i, j, k); x[q][c] = cj * r[q][c]; }
        for (int sweep = 1; sweep < kMgCoarsestSweeps; ++sweep) {
            for (int tid = 0; tid < T; ++tid) mg_smooth<Real>(d[q], s, r[q].data(), x[q].data(), t[q].data(), tid, T);
            x[q].swap(t[q]);
        }
    }
    double dot = 0;
    for (int l = L - 2; l >= 0; --l) {
        s /= 2;
        dot = 0;
        for (int tid = 0; tid < T; ++tid) dot += mg3_up<Real>(d[l], d[l + 1], s, r[l].data(), x[l + 1].data(), x[l].data(), tid, T);
    }
    if (L == 1) { dot = 0; for (size_t c = 0; c < x[0].size(); ++c) dot += (double)r[0][c] * (double)x[0][c]; }
    f = fopen(out, "wb");
    fwrite(x[0].data(), sizeof(Real), x[0].size(), f);
    fclose(f);
    printf("levels %d rz %.17g\n", L, dot);
    return 0;
}

int main(int argc, char** argv) {
    if (argc != 10) { fprintf(stderr, "usage: %s X Y Z j0 j1 prec threads in.bin out.bin\n", argv[0]); return 1; }
    MgDims d{atoi(argv[1]), atoi(argv[2]), atoi(argv[3]), atoi(argv[4]), atoi(argv[5])};
    const int prec = atoi(argv[6]), T = atoi(argv[7]);
    return prec == 64 ? run<double>(d, T, argv[8], argv[9]) : run<float>(d, T, argv[8], argv[9]);
}
