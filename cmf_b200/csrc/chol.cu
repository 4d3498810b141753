// Kernel (c): batched shared-memory Cholesky -> log det G, the canonical metric penalties
// sum_{i!=j}|G_ij| / sum_k|G_kk|, and the adjoint  g_G = g_logdet * G^{-1} + g_l1 * sign(G) (.) mask.
//
// One sample is owned by TPS = next_pow2(d) threads (4..128); a 128-thread CTA carries 128/TPS
// samples, so the 2-D / tabular configs (d = 2..10, B = 1000..5000) pack 16..32 samples per CTA and
// the image configs (d = 64, 128) use 2 / 1.  The factor lives in shared memory with a padded row
// (d+1) so column-of-threads accesses are bank-conflict free.  HBM traffic: G is read once and L
// written once in the forward (8 d^2 B bytes), L and G read once and g_G written once in the
// backward (12 d^2 B bytes) -- negligible next to kernel (a); the kernel is latency-bound by the d
// sequential pivots, which is why several samples share a CTA and many CTAs share an SM.
//
// Left-looking column Cholesky: at step k every thread i >= k forms
//     dot_i = sum_{m<k} L[i][m] L[k][m]      and (redundantly)  p = sum_{m<k} L[k][m]^2,
// so the pivot sqrt(G[k][k] - p) needs no extra barrier: one __syncthreads per column.
#include <math.h>

#include "common.cuh"

namespace cmf {

constexpr int kCholThreads = 128;
constexpr int kCholMaxD = 128;

__device__ __forceinline__ float group_sum(float v, float* red, int tps, int lane) {
    // sum over the tps threads of one sample; red points at this sample's tps-float scratch
    red[lane] = v;
    __syncthreads();
    float s = 0.f;
    for (int i = 0; i < tps; ++i) s += red[i];
    __syncthreads();
    return s;
}

__global__ void __launch_bounds__(kCholThreads)
chol_logdet_fwd_kernel(int B, int d, int tps, const float* __restrict__ G, float* __restrict__ L,
                       float* __restrict__ logdet, float* __restrict__ l1_off, float* __restrict__ l1_diag,
                       int32_t* __restrict__ info) {
    extern __shared__ float smem[];
    const int spb = kCholThreads / tps;
    const int slot = threadIdx.x / tps, lane = threadIdx.x % tps;
    const int b = blockIdx.x * spb + slot;
    const bool live = b < B;
    const int ld = d + 1;
    float* Ls = smem + (size_t)slot * d * ld;
    float* red = smem + (size_t)spb * d * ld + slot * tps;

    float off = 0.f, dia = 0.f;
    if (live) {
        const float* Gb = G + (size_t)b * d * d;
        for (int e = lane; e < d * d; e += tps) {
            const int i = e / d, j = e - i * d;
            const float g = __ldg(Gb + e);
            Ls[i * ld + j] = g;
            if (i == j) dia += fabsf(g); else off += fabsf(g);
        }
    }
    __syncthreads();

    // Every thread i >= k derives the same pivot from the same operands in the same order, so a
    // failure is seen by all of them at once; threads i < k are already idle.  The per-sample flag
    // only carries the verdict to lane 0 for the final write.
    int* failed = reinterpret_cast<int*>(smem + (size_t)spb * d * ld + kCholThreads) + slot;
    if (lane == 0) *failed = 0;
    bool ok = true;
    float mylog = 0.f, mypiv = 0.f;
    const int i = lane;
    for (int k = 0; k < d; ++k) {
        const bool active = live && ok && i >= k && i < d;
        if (active) {
            const float* ri = Ls + i * ld;
            const float* rk = Ls + k * ld;
            float dot = 0.f, p = 0.f;
            for (int m = 0; m < k; ++m) {
                const float lk = rk[m];
                dot = fmaf(ri[m], lk, dot);
                p = fmaf(lk, lk, p);
            }
            const float piv2 = rk[k] - p;      // rk[k] still holds G[k][k]: see the commit below
            if (!(piv2 > 0.f)) {               // not positive definite (or NaN): torch raises here
                ok = false;
                if (i == k) { *failed = 1; atomicMax(info, b + 1); }
            } else {
                const float piv = sqrtf(piv2);
                if (i == k) { mypiv = piv; mylog = logf(piv); }
                else Ls[i * ld + k] = (ri[k] - dot) / piv;
            }
        }
        __syncthreads();
        // the diagonal is committed only now: during step k the other threads still read G[k][k]
        // from this slot, and no later step reads L[k][k].
        if (active && ok && i == k) Ls[k * ld + k] = mypiv;
    }
    const bool sample_ok = live ? (*failed == 0) : true;

    const float ld_sum = group_sum(mylog, red, tps, lane);
    const float off_sum = group_sum(off, red, tps, lane);
    const float dia_sum = group_sum(dia, red, tps, lane);
    if (live) {
        if (lane == 0) {
            logdet[b] = sample_ok ? 2.f * ld_sum : nanf("");
            l1_off[b] = off_sum;
            l1_diag[b] = dia_sum;
        }
        float* Lb = L + (size_t)b * d * d;
        for (int e = lane; e < d * d; e += tps) {
            const int r = e / d, c = e - r * d;
            Lb[e] = c <= r ? Ls[r * ld + c] : 0.f;
        }
    }
}

__global__ void __launch_bounds__(kCholThreads)
chol_logdet_bwd_kernel(int B, int d, int tps, const float* __restrict__ L, const float* __restrict__ G,
                       const float* __restrict__ g_logdet, const float* __restrict__ g_l1, int penalty,
                       float* __restrict__ gG) {
    extern __shared__ float smem[];
    const int spb = kCholThreads / tps;
    const int slot = threadIdx.x / tps, lane = threadIdx.x % tps;
    const int b = blockIdx.x * spb + slot;
    const bool live = b < B;
    const int ld = d + 1;
    float* Ls = smem + (size_t)slot * 2 * d * ld;
    float* Ws = Ls + (size_t)d * ld;   // W = L^{-1}, lower triangular

    if (live) {
        const float* Lb = L + (size_t)b * d * d;
        for (int e = lane; e < d * d; e += tps) {
            const int i = e / d, j = e - i * d;
            Ls[i * ld + j] = __ldg(Lb + e);
        }
    }
    __syncthreads();
    // column j of W by forward substitution (thread-private, no barriers)
    if (live && lane < d) {
        const int j = lane;
        for (int i = 0; i < j; ++i) Ws[i * ld + j] = 0.f;
        Ws[j * ld + j] = 1.f / Ls[j * ld + j];
        for (int i = j + 1; i < d; ++i) {
            const float* ri = Ls + i * ld;
            float acc = 0.f;
            for (int m = j; m < i; ++m) acc = fmaf(ri[m], Ws[m * ld + j], acc);
            Ws[i * ld + j] = -acc / ri[i];
        }
    }
    __syncthreads();
    // G^{-1}[i][j] = sum_{m >= max(i,j)} W[m][i] W[m][j];  thread j owns column j
    if (live && lane < d) {
        const int j = lane;
        const float gl = __ldg(g_logdet + b);
        const float g1 = (penalty != CMF_PENALTY_NONE && g_l1) ? __ldg(g_l1 + b) : 0.f;
        const float* Gb = G + (size_t)b * d * d;
        float* Ob = gG + (size_t)b * d * d;
        for (int i = 0; i < d; ++i) {
            float acc = 0.f;
            for (int m = max(i, j); m < d; ++m) acc = fmaf(Ws[m * ld + i], Ws[m * ld + j], acc);
            float out = gl * acc;
            const bool sel = penalty == CMF_PENALTY_OFFDIAG ? (i != j) : (penalty == CMF_PENALTY_DIAG ? (i == j) : false);
            if (sel) {
                const float g = __ldg(Gb + (size_t)i * d + j);
                out += g1 * (g > 0.f ? 1.f : (g < 0.f ? -1.f : 0.f));
            }
            Ob[(size_t)i * d + j] = out;
        }
    }
}

static int pick_tps(int d) {
    int t = 4;
    while (t < d) t <<= 1;
    return t;
}

}  // namespace cmf

using namespace cmf;

// 128 < d <= 1024: panel-blocked kernels working out of global memory (chol_big.cu)
constexpr int kCholBigMaxD = 1024;
int chol_big_fwd_launch(int B, int d, const float* G, float* L, float* logdet, float* l1_off, float* l1_diag,
                        int32_t* info, cudaStream_t stream);
int chol_big_bwd_launch(int B, int d, const float* L, const float* G, const float* g_logdet, const float* g_l1,
                        int penalty, float* g_G, cudaStream_t stream);

extern "C" int cmf_chol_logdet_fwd(int B, int d, const float* G, float* L, float* logdet,
                                   float* l1_off, float* l1_diag, int32_t* info, cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && d >= 1 && d <= kCholBigMaxD, "chol_fwd: need 1 <= d <= %d (d=%d)", kCholBigMaxD, d);
    if (B == 0) return 0;
    CMF_REQUIRE(G && L && logdet && l1_off && l1_diag && info, "chol_fwd: null pointer");
    if (d > kCholMaxD) return chol_big_fwd_launch(B, d, G, L, logdet, l1_off, l1_diag, info, (cudaStream_t)stream);
    const int tps = pick_tps(d), spb = kCholThreads / tps;
    const size_t smem = ((size_t)spb * d * (d + 1) + kCholThreads + spb) * sizeof(float);
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(chol_logdet_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) { set_error("chol_fwd attr: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set = true;
    }
    chol_logdet_fwd_kernel<<<(unsigned)ceil_div(B, spb), kCholThreads, smem, (cudaStream_t)stream>>>(
        B, d, tps, G, L, logdet, l1_off, l1_diag, info);
    CMF_CHECK_LAUNCH();
    return 0;
}

extern "C" int cmf_chol_logdet_bwd(int B, int d, const float* L, const float* G,
                                   const float* g_logdet, const float* g_l1, int penalty,
                                   float* g_G, cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && d >= 1 && d <= kCholBigMaxD, "chol_bwd: need 1 <= d <= %d (d=%d)", kCholBigMaxD, d);
    CMF_REQUIRE(penalty >= 0 && penalty <= 2, "chol_bwd: bad penalty %d", penalty);
    if (B == 0) return 0;
    CMF_REQUIRE(L && G && g_logdet && g_G, "chol_bwd: null pointer");
    if (d > kCholMaxD) return chol_big_bwd_launch(B, d, L, G, g_logdet, g_l1, penalty, g_G, (cudaStream_t)stream);
    const int tps = pick_tps(d), spb = kCholThreads / tps;
    const size_t smem = (size_t)spb * 2 * d * (d + 1) * sizeof(float);
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(chol_logdet_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) { set_error("chol_bwd attr: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set = true;
    }
    chol_logdet_bwd_kernel<<<(unsigned)ceil_div(B, spb), kCholThreads, smem, (cudaStream_t)stream>>>(
        B, d, tps, L, G, g_logdet, g_l1, penalty, g_G);
    CMF_CHECK_LAUNCH();
    return 0;
}
