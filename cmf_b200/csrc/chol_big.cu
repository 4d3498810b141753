// Kernel (c) for 128 < d <= 1024 (the CIFAR-shape latent sweep, BASELINE configs[4]): the factor
// no longer fits shared memory next to its inverse, so the matrix stays in global memory (L2
// resident: d = 512 is 1 MB per sample) and is processed in 32-wide panels staged in shared memory.
//
//   forward  : one CTA per sample, right-looking blocked Cholesky
//                 diag block (32x32, one warp, in smem) -> panel solve (thread per row, registers)
//                 -> trailing SYRK update (64x64 block tiles, 4x4 register tiles, panel in smem)
//              + log det, both metric penalties, the not-PD flag; same contract as chol.cu.
//   backward : one CTA per (sample, 32-column block c): X = G^{-1}[:, c] by the two triangular
//              solves  L Y = E_c,  L^T X = Y  restricted to rows >= c0 (G^{-1} is symmetric; the
//              mirror is written from the same registers), then
//                 g_G = g_logdet * G^{-1} + g_l1 * sign(G) (.) mask.
// FLOPs: d^3/3 forward, ~2 d^3/3 backward -- 45 / 90 MFLOP per sample at d = 512, far below the
// tangent pushes that produce G; the design goal is parity and no host round trips.
#include <math.h>

#include "common.cuh"

namespace cmf {

constexpr int kNb = 32;
constexpr int kBigThreads = 256;
constexpr int kPld = kNb + 1;     // padded panel row

// factor the nb x nb block D (row stride kPld) in place with one warp (lane = row); returns false
// (to every lane) at the first non-positive pivot.  *logsum accumulates log(pivot) on lane k.
__device__ __forceinline__ bool warp_chol32(float* D, int nb, int lane, float* logsum) {
    bool ok = true;
    for (int k = 0; k < nb; ++k) {
        float dot = 0.f, p = 0.f;
        if (lane >= k && lane < nb) {
            const float* ri = D + lane * kPld;
            const float* rk = D + k * kPld;
            for (int m = 0; m < k; ++m) {
                const float lk = rk[m];
                dot = fmaf(ri[m], lk, dot);
                p = fmaf(lk, lk, p);
            }
        }
        const float piv2 = __shfl_sync(0xffffffffu, D[k * kPld + k] - p, k);   // lane k's value
        if (!(piv2 > 0.f)) { ok = false; break; }
        const float piv = sqrtf(piv2);
        __syncwarp();
        if (lane == k) { D[k * kPld + k] = piv; *logsum += logf(piv); }
        else if (lane > k && lane < nb) D[lane * kPld + k] = (D[lane * kPld + k] - dot) / piv;
        __syncwarp();
    }
    return ok;
}

__global__ void __launch_bounds__(kBigThreads)
chol_big_fwd_kernel(int d, const float* __restrict__ G, float* __restrict__ L,
                    float* __restrict__ logdet, float* __restrict__ l1_off, float* __restrict__ l1_diag,
                    int32_t* __restrict__ info) {
    extern __shared__ float smem[];
    float* Dk = smem;                      // [32][33]
    float* Pn = smem + kNb * kPld;         // [(d - 32)][33] panel below the diagonal block
    __shared__ float red[kBigThreads / 32][2];
    __shared__ int failed;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* Gb = G + (size_t)b * d * d;
    float* Lb = L + (size_t)b * d * d;

    // ---- penalties + working copy (lower triangle; upper = 0)
    float off = 0.f, dia = 0.f;
    for (int e = tid; e < d * d; e += kBigThreads) {
        const int i = e / d, j = e - i * d;
        const float g = __ldg(Gb + e);
        if (i == j) dia += fabsf(g); else off += fabsf(g);
        Lb[e] = j <= i ? g : 0.f;
    }
    off = warp_sum(off); dia = warp_sum(dia);
    if (lane == 0) { red[warp][0] = off; red[warp][1] = dia; }
    if (tid == 0) failed = 0;
    __syncthreads();
    if (tid == 0) {
        float o = 0.f, q = 0.f;
        for (int w = 0; w < kBigThreads / 32; ++w) { o += red[w][0]; q += red[w][1]; }
        l1_off[b] = o; l1_diag[b] = q;
    }

    float logsum = 0.f;
    for (int k0 = 0; k0 < d; k0 += kNb) {
        const int nb = min(kNb, d - k0);
        const int k1 = k0 + nb, m = d - k1;               // trailing size
        for (int e = tid; e < nb * nb; e += kBigThreads) {
            const int i = e / nb, j = e - i * nb;
            Dk[i * kPld + j] = Lb[(size_t)(k0 + i) * d + k0 + j];
        }
        __syncthreads();
        if (warp == 0) {
            const bool ok = warp_chol32(Dk, nb, lane, &logsum);
            if (!ok && lane == 0) failed = 1;
        }
        __syncthreads();
        if (failed) break;
        for (int e = tid; e < nb * nb; e += kBigThreads) {
            const int i = e / nb, j = e - i * nb;
            if (j <= i) Lb[(size_t)(k0 + i) * d + k0 + j] = Dk[i * kPld + j];
        }
        // ---- panel solve: row r of the panel, L[r, k0:k1] = A[r, k0:k1] * Lkk^{-T}
        for (int r = k1 + tid; r < d; r += kBigThreads) {
            float a[kNb];
            float* row = Lb + (size_t)r * d + k0;
#pragma unroll
            for (int j = 0; j < kNb; ++j) a[j] = j < nb ? row[j] : 0.f;
#pragma unroll
            for (int j = 0; j < kNb; ++j) {
                if (j < nb) {
                    float s = a[j];
#pragma unroll
                    for (int q = 0; q < j; ++q) s = fmaf(-a[q], Dk[j * kPld + q], s);
                    a[j] = s / Dk[j * kPld + j];
                }
            }
#pragma unroll
            for (int j = 0; j < kNb; ++j) {
                if (j < nb) row[j] = a[j];
                Pn[(r - k1) * kPld + j] = a[j];
            }
        }
        __syncthreads();
        // ---- trailing update A22 -= P P^T on the lower triangle, 64x64 block tiles
        const int nbt = (m + 63) / 64;
        const int ty = tid >> 4, tx = tid & 15;
        for (int bi = 0; bi < nbt; ++bi) {
            for (int bj = 0; bj <= bi; ++bj) {
                float acc[4][4] = {};
                const int i0 = bi * 64 + ty, j0 = bj * 64 + tx;
                for (int t = 0; t < nb; ++t) {
                    float av[4], bv[4];
#pragma unroll
                    for (int x = 0; x < 4; ++x) {
                        const int i = i0 + 16 * x, j = j0 + 16 * x;
                        av[x] = i < m ? Pn[i * kPld + t] : 0.f;
                        bv[x] = j < m ? Pn[j * kPld + t] : 0.f;
                    }
#pragma unroll
                    for (int x = 0; x < 4; ++x)
#pragma unroll
                        for (int y = 0; y < 4; ++y) acc[x][y] = fmaf(av[x], bv[y], acc[x][y]);
                }
#pragma unroll
                for (int x = 0; x < 4; ++x)
#pragma unroll
                    for (int y = 0; y < 4; ++y) {
                        const int i = i0 + 16 * x, j = j0 + 16 * y;
                        if (i < m && j <= i) Lb[(size_t)(k1 + i) * d + k1 + j] -= acc[x][y];
                    }
            }
        }
        __syncthreads();
    }
    if (warp == 0) {
        logsum = warp_sum(logsum);
        if (lane == 0) {
            const bool bad = failed != 0;
            logdet[b] = bad ? nanf("") : 2.f * logsum;
            if (bad) atomicMax(info, b + 1);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// backward: CTA (b, cb) -> columns [c0, c0+32) of G^{-1}, rows >= c0
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBigThreads)
chol_big_bwd_kernel(int d, const float* __restrict__ L, const float* __restrict__ G,
                    const float* __restrict__ g_logdet, const float* __restrict__ g_l1, int penalty,
                    float* __restrict__ gG) {
    extern __shared__ float smem[];
    float* Lt = smem;                       // [32][33] off-diagonal tile of L
    float* Ld = smem + kNb * kPld;          // [32][33] diagonal block of L
    float* Ys = smem + 2 * kNb * kPld;      // [(d - c0)][33]  Y, then X in place
    const int b = blockIdx.x, c0 = blockIdx.y * kNb;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* Lb = L + (size_t)b * d * d;
    const int mrows = d - c0;
    const int nrb = (mrows + kNb - 1) / kNb;

    // ---- forward substitution  L Y = E_c  (rows c0 .. d-1; rows above c0 of Y are zero)
    for (int rb = 0; rb < nrb; ++rb) {
        const int r0 = c0 + rb * kNb, nr = min(kNb, d - r0);
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int kt = 0; kt < rb; ++kt) {
            const int k0 = c0 + kt * kNb;
            for (int e = tid; e < kNb * kNb; e += kBigThreads) {
                const int i = e >> 5, kk = e & 31;
                Lt[i * kPld + kk] = i < nr ? __ldg(Lb + (size_t)(r0 + i) * d + k0 + kk) : 0.f;
            }
            __syncthreads();
#pragma unroll 8
            for (int kk = 0; kk < kNb; ++kk) {
                const float y = Ys[(kt * kNb + kk) * kPld + lane];
#pragma unroll
                for (int x = 0; x < 4; ++x) acc[x] = fmaf(Lt[(warp * 4 + x) * kPld + kk], y, acc[x]);
            }
            __syncthreads();
        }
        for (int e = tid; e < kNb * kNb; e += kBigThreads) {
            const int i = e >> 5, j = e & 31;
            Ld[i * kPld + j] = (i < nr && j < nr) ? __ldg(Lb + (size_t)(r0 + i) * d + r0 + j) : 0.f;
        }
#pragma unroll
        for (int x = 0; x < 4; ++x) {
            const int i = warp * 4 + x;
            if (i < nr) Ys[(rb * kNb + i) * kPld + lane] = ((rb == 0 && i == lane) ? 1.f : 0.f) - acc[x];
        }
        __syncthreads();
        if (warp == 0) {
            for (int i = 0; i < nr; ++i) {
                float y = Ys[(rb * kNb + i) * kPld + lane];
                for (int m = 0; m < i; ++m) y = fmaf(-Ld[i * kPld + m], Ys[(rb * kNb + m) * kPld + lane], y);
                Ys[(rb * kNb + i) * kPld + lane] = y / Ld[i * kPld + i];
            }
        }
        __syncthreads();
    }
    // ---- backward substitution  L^T X = Y  (X overwrites Y), last row block first
    for (int rb = nrb - 1; rb >= 0; --rb) {
        const int r0 = c0 + rb * kNb, nr = min(kNb, d - r0);
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        for (int kt = rb + 1; kt < nrb; ++kt) {
            const int k0 = c0 + kt * kNb, nkk = min(kNb, d - k0);
            for (int e = tid; e < kNb * kNb; e += kBigThreads) {
                const int kk = e >> 5, i = e & 31;      // Lt[kk][i] = L[k0+kk][r0+i]
                Lt[kk * kPld + i] = (kk < nkk && i < nr) ? __ldg(Lb + (size_t)(k0 + kk) * d + r0 + i) : 0.f;
            }
            __syncthreads();
            for (int kk = 0; kk < nkk; ++kk) {
                const float xk = Ys[(kt * kNb + kk) * kPld + lane];
#pragma unroll
                for (int x = 0; x < 4; ++x) acc[x] = fmaf(Lt[kk * kPld + warp * 4 + x], xk, acc[x]);
            }
            __syncthreads();
        }
        for (int e = tid; e < kNb * kNb; e += kBigThreads) {
            const int i = e >> 5, j = e & 31;
            Ld[i * kPld + j] = (i < nr && j < nr) ? __ldg(Lb + (size_t)(r0 + i) * d + r0 + j) : 0.f;
        }
#pragma unroll
        for (int x = 0; x < 4; ++x) {
            const int i = warp * 4 + x;
            if (i < nr) Ys[(rb * kNb + i) * kPld + lane] -= acc[x];
        }
        __syncthreads();
        if (warp == 0) {
            for (int i = nr - 1; i >= 0; --i) {
                float xv = Ys[(rb * kNb + i) * kPld + lane];
                for (int m = i + 1; m < nr; ++m) xv = fmaf(-Ld[m * kPld + i], Ys[(rb * kNb + m) * kPld + lane], xv);
                Ys[(rb * kNb + i) * kPld + lane] = xv / Ld[i * kPld + i];
            }
        }
        __syncthreads();
    }
    // ---- g_G = g_logdet * G^{-1} + g_l1 * sign(G) (.) mask; direct block rows then the mirror
    const int nc = min(kNb, d - c0);
    const float gl = __ldg(g_logdet + b);
    const float g1 = (penalty != CMF_PENALTY_NONE && g_l1) ? __ldg(g_l1 + b) : 0.f;
    const float* Gb = G + (size_t)b * d * d;
    float* Ob = gG + (size_t)b * d * d;
    auto emit = [&](int i, int j, float ginv) {
        float out = gl * ginv;
        const bool sel = penalty == CMF_PENALTY_OFFDIAG ? (i != j) : (penalty == CMF_PENALTY_DIAG ? (i == j) : false);
        if (sel) {
            const float g = __ldg(Gb + (size_t)i * d + j);
            out += g1 * (g > 0.f ? 1.f : (g < 0.f ? -1.f : 0.f));
        }
        Ob[(size_t)i * d + j] = out;
    };
    for (int e = tid; e < mrows * kNb; e += kBigThreads) {      // rows r >= c0, lanes along the column
        const int rr = e >> 5, c = e & 31;
        if (c < nc) emit(c0 + rr, c0 + c, Ys[rr * kPld + c]);
    }
    const int mr2 = mrows - nc;                                  // rows below the diagonal block
    for (int e = tid; e < mr2 * nc; e += kBigThreads) {          // mirror: lanes along r (coalesced)
        const int c = e / mr2, rr = nc + (e - c * mr2);
        emit(c0 + c, c0 + rr, Ys[rr * kPld + c]);
    }
}

}  // namespace cmf

using namespace cmf;

int chol_big_fwd_launch(int B, int d, const float* G, float* L, float* logdet, float* l1_off, float* l1_diag,
                        int32_t* info, cudaStream_t stream) {
    const size_t smem = ((size_t)kNb * kPld + (size_t)max(d - kNb, 1) * kPld) * sizeof(float);
    static DeviceOnce attr_set;
    if (!attr_set.done()) {
        cudaError_t e = cudaFuncSetAttribute(chol_big_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) { set_error("chol_big_fwd attr: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set.mark();
    }
    chol_big_fwd_kernel<<<(unsigned)B, kBigThreads, smem, stream>>>(d, G, L, logdet, l1_off, l1_diag, info);
    CMF_CHECK_LAUNCH();
    return 0;
}

int chol_big_bwd_launch(int B, int d, const float* L, const float* G, const float* g_logdet, const float* g_l1,
                        int penalty, float* g_G, cudaStream_t stream) {
    const size_t smem = ((size_t)2 * kNb * kPld + (size_t)d * kPld) * sizeof(float);
    static DeviceOnce attr_set;
    if (!attr_set.done()) {
        cudaError_t e = cudaFuncSetAttribute(chol_big_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) { set_error("chol_big_bwd attr: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set.mark();
    }
    dim3 grid((unsigned)B, (unsigned)ceil_div(d, kNb));
    chol_big_bwd_kernel<<<grid, kBigThreads, smem, stream>>>(d, L, G, g_logdet, g_l1, penalty, g_G);
    CMF_CHECK_LAUNCH();
    return 0;
}
