// Index maps folded into addressing: one gather primitive covers the tail zero-pad + inverse
// permutation (non_square.py:397-410), the split zero-pad (split.py:50-52), squeeze/unsqueeze
// (reshaping.py:102-114), checkerboard masking of conditioner inputs (acl.py:54) and the
// re-partitioning of compact halves between layer groups.  Its adjoint is the same primitive with
// the inverse index (the maps are partial permutations).
//   out[r, i] = idx[i] < 0 ? 0 : (idx[i] < n0 ? in0[r, idx[i]] : in1[r, idx[i] - n0])
// Writes are dense and coalesced (thread <-> output element); reads are coalesced whenever idx is
// locally monotone, which holds for every map the layer program builds.
#include "common.cuh"

namespace cmf {

constexpr int kGatherThreads = 256;

__global__ void __launch_bounds__(kGatherThreads)
gather_rows_kernel(long long rows, int n0, int n1, int n_out,
                   const float* __restrict__ in0, const float* __restrict__ in1,
                   const int32_t* __restrict__ idx, float* __restrict__ out) {
    const long long total = rows * n_out;
    const long long stride = (long long)gridDim.x * kGatherThreads;
    for (long long t = (long long)blockIdx.x * kGatherThreads + threadIdx.x; t < total; t += stride) {
        const long long r = t / n_out;
        const int i = (int)(t - r * n_out);
        const int s = __ldg(idx + i);
        float v = 0.f;
        if (s >= 0) v = s < n0 ? __ldcs(in0 + r * n0 + s) : __ldcs(in1 + r * n1 + (s - n0));
        out[t] = v;
    }
}

}  // namespace cmf

extern "C" int cmf_gather_rows(long long rows, int n0, int n1, int n_out,
                               const float* in0, const float* in1, const int32_t* idx,
                               float* out, cmf_stream_t stream) {
    using namespace cmf;
    CMF_REQUIRE(rows >= 0 && n0 >= 0 && n1 >= 0 && n_out >= 0, "gather_rows: negative size");
    if (rows == 0 || n_out == 0) return 0;
    CMF_REQUIRE(idx && out, "gather_rows: null idx/out");
    CMF_REQUIRE((n0 == 0 || in0) && (n1 == 0 || in1), "gather_rows: null input with non-zero width");
    const long long total = rows * n_out;
    long long blocks = ceil_div(total, kGatherThreads);
    const long long cap = 8LL * kNumSMs * 4;  // grid-stride beyond ~32 resident blocks per SM
    if (blocks > cap) blocks = cap;
    gather_rows_kernel<<<(unsigned)blocks, kGatherThreads, 0, (cudaStream_t)stream>>>(
        rows, n0, n1, n_out, in0, in1, idx, out);
    CMF_CHECK_LAUNCH();
    return 0;
}
