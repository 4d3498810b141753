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
constexpr int kCholWideMinD = 32;     // from here up: one sample per CTA (chol_wide_*)

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


// ---------------------------------------------------------------------------------------------
// 32 <= d <= 128 (the image configs: d = 64 MNIST-shape, 64..128 CIFAR-shape): one sample per CTA.
// Same left-looking algorithm and the same contract as above, restructured for latency: rows are
// padded to ld floats with ld % 4 == 0 and ld / 4 odd, so a thread reads ITS row (and, broadcast,
// the pivot row) with conflict-free 128-bit loads and the dot products run on four independent FMA
// chains -- the scalar version above is bound by one LDS latency per FMA.  B = 512, d = 64:
// forward 79 -> see profiles; backward: W = L^-1 by columns (128-bit loads of the L row), then
// G^-1 = W^T W as a 4x4 register-tiled product over all 256 threads instead of one column per thread.
// ---------------------------------------------------------------------------------------------
constexpr int kWideFwdThreads = 256;
constexpr int kWideBwdThreads = 256;

__host__ __device__ inline int wide_ld(int d) {
    int ld = (d + 3) & ~3;
    if (((ld >> 2) & 1) == 0) ld += 4;
    return ld;
}

__device__ __forceinline__ float block_sum(float v, float* red) {     // red: >= 8 floats; all threads call
    v = warp_sum(v);
    const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[w] = v;
    __syncthreads();
    float s = 0.f;
    for (int i = 0; i < nw; ++i) s += red[i];
    return s;
}

// tpr = threads per row (a power of two <= 8, d * tpr <= 256): the tpr lanes of a row split the
// 128-bit chunks of the two dot products and combine them with xor shuffles, so a column step is
// one short chain (load, <= 2 FMAs deep per 16 columns, two shuffles, rsqrt, multiply, store,
// barrier) instead of k dependent iterations.  L grows in a zero-initialised buffer of its own:
// chunks may then run past column k (the tail multiplies zeros) and need no scalar remainder.
__device__ __forceinline__ float rsqrt_nr(float x) {       // 1 / sqrt(x) to ~1 ulp
    const float r = rsqrtf(x);
    return r * fmaf(-0.5f * x * r, r, 1.5f);
}

__global__ void __launch_bounds__(kWideFwdThreads)
chol_wide_fwd_kernel(int d, int ld, int tpr, int vec, const float* __restrict__ G, float* __restrict__ L,
                     float* __restrict__ logdet, float* __restrict__ l1_off, float* __restrict__ l1_diag,
                     int32_t* __restrict__ info) {
    extern __shared__ __align__(16) float smem[];
    float* Gs = smem;                          // [d][ld]  the input
    float* Lz = smem + (size_t)d * ld;         // [d][ld]  the factor, zero where not (yet) written
    float* red = Lz + (size_t)d * ld;          // [8]
    __shared__ int failed;
    const int b = blockIdx.x, t = threadIdx.x;
    const int i = t / tpr, sub = t - i * tpr;
    const float* Gb = G + (size_t)b * d * d;
    float off = 0.f, dia = 0.f;
    if (t == 0) failed = 0;
    for (int e = t; e < d * ld; e += kWideFwdThreads) Lz[e] = 0.f;
    if (vec) {
        const int d4 = d >> 2;
        for (int e = t; e < d * d4; e += kWideFwdThreads) {
            const int r = e / d4, c = (e - r * d4) << 2;
            const float4 g = ldg4(Gb + (size_t)r * d + c);
            *reinterpret_cast<float4*>(Gs + r * ld + c) = g;
            const float a[4] = {fabsf(g.x), fabsf(g.y), fabsf(g.z), fabsf(g.w)};
#pragma unroll
            for (int q = 0; q < 4; ++q) { if (c + q == r) dia += a[q]; else off += a[q]; }
        }
    } else {
        for (int e = t; e < d * d; e += kWideFwdThreads) {
            const int r = e / d, c = e - r * d;
            const float g = __ldg(Gb + e);
            Gs[r * ld + c] = g;
            if (r == c) dia += fabsf(g); else off += fabsf(g);
        }
    }
    __syncthreads();

    bool ok = true;
    float mylog = 0.f, mypiv = 0.f;
    const float4* ri4 = reinterpret_cast<const float4*>(Lz + (size_t)(i < d ? i : 0) * ld);
    for (int k = 0; k < d; ++k) {
        const bool active = ok && i >= k && i < d;
        float dot = 0.f, p = 0.f;
        if (active) {
            const float4* rk4 = reinterpret_cast<const float4*>(Lz + (size_t)k * ld);
            const int nq = (k + 3) >> 2;
            float d1 = 0.f, p1 = 0.f;
            int q = sub;
            for (; q + tpr < nq; q += 2 * tpr) {
                const float4 a = ri4[q], c = rk4[q], a2 = ri4[q + tpr], c2 = rk4[q + tpr];
                dot = fmaf(a.x, c.x, dot); dot = fmaf(a.y, c.y, dot); dot = fmaf(a.z, c.z, dot); dot = fmaf(a.w, c.w, dot);
                p = fmaf(c.x, c.x, p); p = fmaf(c.y, c.y, p); p = fmaf(c.z, c.z, p); p = fmaf(c.w, c.w, p);
                d1 = fmaf(a2.x, c2.x, d1); d1 = fmaf(a2.y, c2.y, d1); d1 = fmaf(a2.z, c2.z, d1); d1 = fmaf(a2.w, c2.w, d1);
                p1 = fmaf(c2.x, c2.x, p1); p1 = fmaf(c2.y, c2.y, p1); p1 = fmaf(c2.z, c2.z, p1); p1 = fmaf(c2.w, c2.w, p1);
            }
            if (q < nq) {
                const float4 a = ri4[q], c = rk4[q];
                dot = fmaf(a.x, c.x, dot); dot = fmaf(a.y, c.y, dot); dot = fmaf(a.z, c.z, dot); dot = fmaf(a.w, c.w, dot);
                p = fmaf(c.x, c.x, p); p = fmaf(c.y, c.y, p); p = fmaf(c.z, c.z, p); p = fmaf(c.w, c.w, p);
            }
            dot += d1; p += p1;
        }
        for (int o = tpr >> 1; o > 0; o >>= 1) {            // the tpr lanes of a row are adjacent in one warp
            dot += __shfl_xor_sync(0xffffffffu, dot, o);
            p += __shfl_xor_sync(0xffffffffu, p, o);
        }
        if (active) {
            const float piv2 = Gs[k * ld + k] - p;
            if (!(piv2 > 0.f)) {               // not positive definite (or NaN): torch raises here
                ok = false;
                if (i == k && sub == 0) { failed = 1; atomicMax(info, b + 1); }
            } else if (sub == 0) {
                const float rs = rsqrt_nr(piv2);
                if (i == k) { mypiv = piv2 * rs; mylog = 0.5f * logf(piv2); }
                else Lz[i * ld + k] = (Gs[i * ld + k] - dot) * rs;
            }
        }
        __syncthreads();
        // the diagonal is committed only now: chunks read during step k must see L[k][k] as zero
        if (active && ok && i == k && sub == 0) Lz[k * ld + k] = mypiv;
    }
    const float ld_sum = block_sum(mylog, red);
    const float off_sum = block_sum(off, red);
    const float dia_sum = block_sum(dia, red);
    if (t == 0) {
        logdet[b] = failed ? nanf("") : 2.f * ld_sum;
        l1_off[b] = off_sum;
        l1_diag[b] = dia_sum;
    }
    float* Lb = L + (size_t)b * d * d;
    if (vec) {
        const int d4 = d >> 2;
        for (int e = t; e < d * d4; e += kWideFwdThreads) {
            const int r = e / d4, c = (e - r * d4) << 2;
            *reinterpret_cast<float4*>(Lb + (size_t)r * d + c) = *reinterpret_cast<const float4*>(Lz + r * ld + c);
        }
    } else {
        for (int e = t; e < d * d; e += kWideFwdThreads) {
            const int r = e / d, c = e - r * d;
            Lb[e] = Lz[r * ld + c];
        }
    }
}

__global__ void __launch_bounds__(kWideBwdThreads)
chol_wide_bwd_kernel(int d, int ld, int tpc, int vec, int blocked, const float* __restrict__ L, const float* __restrict__ G,
                     const float* __restrict__ g_logdet, const float* __restrict__ g_l1, int penalty,
                     float* __restrict__ gG) {
    extern __shared__ __align__(16) float smem[];
    float* Ls = smem;                          // [d][ld]
    float* Ws = smem + (size_t)d * ld;         // [d][ld]  W = L^{-1} (lower triangular, zero elsewhere incl. padding)
    float* invd = Ws + (size_t)d * ld;         // [d]      1 / L[i][i]
    const int b = blockIdx.x, t = threadIdx.x;
    const float* Lb = L + (size_t)b * d * d;
    if (vec) {
        const int d4 = d >> 2;
        for (int e = t; e < d * d4; e += kWideBwdThreads) {
            const int r = e / d4, c = (e - r * d4) << 2;
            *reinterpret_cast<float4*>(Ls + r * ld + c) = ldg4(Lb + (size_t)r * d + c);
        }
    } else {
        for (int e = t; e < d * d; e += kWideBwdThreads) {
            const int r = e / d, c = e - r * d;
            Ls[r * ld + c] = __ldg(Lb + e);
        }
    }
    for (int e = t; e < d * ld; e += kWideBwdThreads) Ws[e] = 0.f;
    if (t < d) invd[t] = 1.f / __ldg(Lb + (size_t)t * d + t);
    __syncthreads();
    if (blocked) {
        // W = L^{-1} blockwise (d a multiple of 16): forward substitution is a chain of d - 1 dependent row steps (~450
        // cycles each at d = 64: 34 us of latency per sample whatever the batch).  Here only the 16 x 16 diagonal blocks
        // are inverted by substitution (15 steps, all blocks side by side); the blocks below the diagonal follow from
        //     [A 0; B C]^{-1} = [A^{-1} 0; -C^{-1} B A^{-1}  C^{-1}]
        // level by level (block size 16, 32, ...): two small matrix products per level, every element independent.
        float* Ts = invd + d;                          // [d/2][ld] scratch: T = B A^{-1} of the current level
        {
            const int jl = (t & 63) >> 2, sub = t & 3; // 64 threads per diagonal block: 16 columns x 4 lanes
            for (int blk = t >> 6; blk < d / 16; blk += kWideBwdThreads / 64) {
                const int o = 16 * blk, j = o + jl;
                if (sub == 0) Ws[j * ld + j] = invd[j];
                for (int il = 1; il < 16; ++il) {
                    const int i = o + il;
                    float acc = 0.f;
                    if (jl < il) {
                        const float* ri = Ls + i * ld + o;
                        for (int m = jl + sub; m < il; m += 4) acc = fmaf(ri[m], Ws[(o + m) * ld + j], acc);
                    }
                    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
                    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
                    if (jl < il && sub == 0) Ws[i * ld + j] = -acc * invd[i];
                    __syncwarp();
                }
            }
        }
        __syncthreads();
        for (int bs = 16; bs < d; bs <<= 1) {
            // pairs of diagonal blocks of size bs; an odd block at the end (d / bs odd) has no partner at this level and
            // is merged at a later one -- d / 16 need not be a power of two: the pair (A, C) at offset c0 exists while
            // c0 + bs < d, and C is then min(bs, d - c0 - bs) rows (already inverted as a whole by the earlier levels
            // only if it is a full block; so require full pairs: handled by the host, which sets blocked only for d = 16 * 2^k)
            const int npairs = d / (2 * bs);
            const int per = bs * bs;
            for (int e = t; e < npairs * per; e += kWideBwdThreads) {          // T = L_CA W_AA
                const int pr = e / per, rem = e - pr * per;
                const int r = rem / bs, c = rem - r * bs;
                const int c0 = pr * 2 * bs;
                const float* lrow = Ls + (c0 + bs + r) * ld + c0;
                float acc = 0.f;
                for (int m = c; m < bs; ++m) acc = fmaf(lrow[m], Ws[(c0 + m) * ld + c0 + c], acc);   // W_AA is lower triangular
                Ts[(pr * bs + r) * ld + c] = acc;
            }
            __syncthreads();
            for (int e = t; e < npairs * per; e += kWideBwdThreads) {          // W_CA = -W_CC T
                const int pr = e / per, rem = e - pr * per;
                const int r = rem / bs, c = rem - r * bs;
                const int c0 = pr * 2 * bs;
                const float* wrow = Ws + (c0 + bs + r) * ld + c0 + bs;
                float acc = 0.f;
                for (int m = 0; m <= r; ++m) acc = fmaf(wrow[m], Ts[(pr * bs + m) * ld + c], acc);    // W_CC is lower triangular
                Ws[(c0 + bs + r) * ld + c0 + c] = -acc;
            }
            __syncthreads();
        }
    } else
    // column j of W by forward substitution, tpc lanes per column (adjacent lanes of one warp): they
    // split the dot product W[i][j] = -(sum_{m<i} L[i][m] W[m][j]) / L[i][i] by m mod tpc and combine
    // with xor shuffles -- a row step is a ~100-cycle chain instead of i dependent iterations.
    {
        const int j = t / tpc, sub = t - j * tpc;
        const bool col = j < d;
        if (col && sub == 0) Ws[j * ld + j] = invd[j];
        const int m0 = (j / tpc) * tpc + sub;          // first m >= the aligned group of j (W[m][j] = 0 for m < j)
        for (int i = 1; i < d; ++i) {
            float a0 = 0.f, a1 = 0.f;
            if (col && j < i) {
                const float* ri = Ls + i * ld;
                int m = m0;
                for (; m + tpc < i; m += 2 * tpc) {
                    a0 = fmaf(ri[m], Ws[m * ld + j], a0);
                    a1 = fmaf(ri[m + tpc], Ws[(m + tpc) * ld + j], a1);
                }
                if (m < i) a0 = fmaf(ri[m], Ws[m * ld + j], a0);
            }
            float acc = a0 + a1;
            for (int o = tpc >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if (col && j < i && sub == 0) Ws[i * ld + j] = -acc * invd[i];
            __syncwarp();
        }
    }
    __syncthreads();
    // G^{-1} = W^T W in 4x4 register tiles; tile (ty, tx) needs rows m >= max(4 ty, 4 tx) only
    const float gl = __ldg(g_logdet + b);
    const float g1 = (penalty != CMF_PENALTY_NONE && g_l1) ? __ldg(g_l1 + b) : 0.f;
    const float* Gb = G + (size_t)b * d * d;
    float* Ob = gG + (size_t)b * d * d;
    const int nt = (d + 3) >> 2;
    for (int tile = t; tile < nt * nt; tile += kWideBwdThreads) {
        const int ty = tile / nt, tx = tile - ty * nt;
        float acc[4][4] = {};
        for (int m = 4 * max(ty, tx); m < d; ++m) {
            const float4 a = *reinterpret_cast<const float4*>(Ws + m * ld + 4 * ty);
            const float4 c = *reinterpret_cast<const float4*>(Ws + m * ld + 4 * tx);
            const float av[4] = {a.x, a.y, a.z, a.w}, cv[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
            for (int p = 0; p < 4; ++p)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[p][q] = fmaf(av[p], cv[q], acc[p][q]);
        }
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int i = 4 * ty + p;
            if (i >= d) continue;
            float out[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int j = 4 * tx + q;
                out[q] = gl * acc[p][q];
                const bool sel = penalty == CMF_PENALTY_OFFDIAG ? (i != j) : (penalty == CMF_PENALTY_DIAG ? (i == j) : false);
                if (sel && j < d) {
                    const float g = __ldg(Gb + (size_t)i * d + j);
                    out[q] += g1 * (g > 0.f ? 1.f : (g < 0.f ? -1.f : 0.f));
                }
            }
            if (vec) {
                *reinterpret_cast<float4*>(Ob + (size_t)i * d + 4 * tx) = make_float4(out[0], out[1], out[2], out[3]);
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (4 * tx + q < d) Ob[(size_t)i * d + 4 * tx + q] = out[q];
            }
        }
    }
}


// ---------------------------------------------------------------------------------------------
// d == 64 (the MNIST-shape latent): ONE WARP per sample, the matrix in registers.  Lane l owns rows
// l and l + 32 (2 x 64 floats); the loops over k and j are fully unrolled so every register index
// is static.  Right-looking: at step k the pivot is broadcast by one shuffle, the scaled column goes
// through a 64-float shared buffer (two stores, 128-bit broadcast loads) and every lane updates the
// trailing part of its two rows with independent FMAs.  ~80 warp instructions per step instead of
// the ~600 the block-per-sample kernel spends on barriers, shuffles and loop control (measured:
// that kernel is bound by instruction issue of its 8 warps, not by arithmetic).  Entries above the
// diagonal carry garbage that is never read; the pivots decide positive definiteness exactly as
// before.
// ---------------------------------------------------------------------------------------------
constexpr int kWarpD = 64;
constexpr int kWarpCholThreads = 128;       // 4 samples per CTA

__global__ void __launch_bounds__(kWarpCholThreads)
chol_warp64_fwd_kernel(int B, const float* __restrict__ G, float* __restrict__ L,
                       float* __restrict__ logdet, float* __restrict__ l1_off, float* __restrict__ l1_diag,
                       int32_t* __restrict__ info) {
    constexpr int D = kWarpD;
    __shared__ __align__(16) float colbuf[kWarpCholThreads / 32][2][D];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.x * (kWarpCholThreads / 32) + warp;
    if (b >= B) return;                         // warp-uniform; no block-wide barrier below
    const float* Gb = G + (size_t)b * D * D;
    float a0[D], a1[D];
#pragma unroll
    for (int c = 0; c < D / 4; ++c) {
        const float4 u = ldg4(Gb + (size_t)lane * D + 4 * c);
        const float4 v = ldg4(Gb + (size_t)(lane + 32) * D + 4 * c);
        a0[4 * c] = u.x; a0[4 * c + 1] = u.y; a0[4 * c + 2] = u.z; a0[4 * c + 3] = u.w;
        a1[4 * c] = v.x; a1[4 * c + 1] = v.y; a1[4 * c + 2] = v.z; a1[4 * c + 3] = v.w;
    }
    float off = 0.f, dia = 0.f;
#pragma unroll
    for (int j = 0; j < D; ++j) {
        const float x0 = fabsf(a0[j]), x1 = fabsf(a1[j]);
        if (j == lane) dia += x0; else off += x0;
        if (j == lane + 32) dia += x1; else off += x1;
    }
    bool fail = false;
#pragma unroll
    for (int k = 0; k < D; ++k) {
        const float akk = __shfl_sync(0xffffffffu, k < 32 ? a0[k] : a1[k], k & 31);
        if (!(akk > 0.f)) fail = true;          // same value in every lane
        const float rs = rsqrt_nr(akk);
        const float l0 = a0[k] * rs, l1 = a1[k] * rs;
        a0[k] = l0;
        a1[k] = l1;
        if (k + 1 < D) {
            float* col = colbuf[warp][k & 1];
            col[lane] = l0;
            col[lane + 32] = l1;
            __syncwarp();
#pragma unroll
            for (int g = (k + 1) / 4; g < D / 4; ++g) {
                const float4 c = *reinterpret_cast<const float4*>(col + 4 * g);
                const float cv[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int j = 4 * g + e;
                    if (j > k) {
                        a0[j] = fmaf(-l0, cv[e], a0[j]);
                        a1[j] = fmaf(-l1, cv[e], a1[j]);
                    }
                }
            }
        }
    }
    // log det from the diagonal entries this lane owns
    float d0 = 1.f, d1 = 1.f;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        if (j == lane) { d0 = a0[j]; d1 = a1[j + 32]; }
    }
    const float ld_sum = warp_sum(logf(d0) + logf(d1));
    const float off_sum = warp_sum(off), dia_sum = warp_sum(dia);
    if (lane == 0) {
        if (fail) atomicMax(info, b + 1);
        logdet[b] = fail ? nanf("") : 2.f * ld_sum;
        l1_off[b] = off_sum;
        l1_diag[b] = dia_sum;
    }
    float* Lb = L + (size_t)b * D * D;
#pragma unroll
    for (int c = 0; c < D / 4; ++c) {
        float4 u, v;
        u.x = 4 * c + 0 <= lane ? a0[4 * c + 0] : 0.f;
        u.y = 4 * c + 1 <= lane ? a0[4 * c + 1] : 0.f;
        u.z = 4 * c + 2 <= lane ? a0[4 * c + 2] : 0.f;
        u.w = 4 * c + 3 <= lane ? a0[4 * c + 3] : 0.f;
        v.x = 4 * c + 0 <= lane + 32 ? a1[4 * c + 0] : 0.f;
        v.y = 4 * c + 1 <= lane + 32 ? a1[4 * c + 1] : 0.f;
        v.z = 4 * c + 2 <= lane + 32 ? a1[4 * c + 2] : 0.f;
        v.w = 4 * c + 3 <= lane + 32 ? a1[4 * c + 3] : 0.f;
        *reinterpret_cast<float4*>(Lb + (size_t)lane * D + 4 * c) = u;
        *reinterpret_cast<float4*>(Lb + (size_t)(lane + 32) * D + 4 * c) = v;
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
    if (d == kWarpD && aligned16(G) && aligned16(L)) {
        chol_warp64_fwd_kernel<<<(unsigned)ceil_div(B, kWarpCholThreads / 32), kWarpCholThreads, 0, (cudaStream_t)stream>>>(
            B, G, L, logdet, l1_off, l1_diag, info);
        CMF_CHECK_LAUNCH();
        return 0;
    }
    if (d >= kCholWideMinD) {
        const int ld = wide_ld(d);
        const int vec = (d % 4 == 0) && aligned16(G) && aligned16(L);
        const size_t smem = ((size_t)2 * d * ld + 8) * sizeof(float);
        int tpr = 8;
        while (d * tpr > kWideFwdThreads) tpr >>= 1;
        static DeviceOnce wide_attr;                   // the opt-in is per device
        if (!wide_attr.done()) {
            cudaError_t e = cudaFuncSetAttribute(chol_wide_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e != cudaSuccess) { set_error("chol_fwd attr: %s", cudaGetErrorString(e)); return (int)e; }
            wide_attr.mark();
        }
        chol_wide_fwd_kernel<<<(unsigned)B, kWideFwdThreads, smem, (cudaStream_t)stream>>>(d, ld, tpr, vec, G, L, logdet, l1_off, l1_diag, info);
        CMF_CHECK_LAUNCH();
        return 0;
    }
    const int tps = pick_tps(d), spb = kCholThreads / tps;
    const size_t smem = ((size_t)spb * d * (d + 1) + kCholThreads + spb) * sizeof(float);
    static DeviceOnce attr_set;
    if (!attr_set.done()) {
        cudaError_t e = cudaFuncSetAttribute(chol_logdet_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) { set_error("chol_fwd attr: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set.mark();
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
    if (d >= kCholWideMinD) {
        const int ld = wide_ld(d);
        const int vec = (d % 4 == 0) && aligned16(L) && aligned16(g_G);
        // blockwise inverse for d = 16 * 2^k (32, 64, 128): full pairs at every level
        static const bool blocked_off = [] { const char* e = getenv("CMF_CHOL_BWD_BLOCKED"); return e && e[0] == '0'; }();
        const int blocked = !blocked_off && d % 16 == 0 && ((d / 16) & (d / 16 - 1)) == 0;
        const size_t smem = ((size_t)2 * d * ld + d + (blocked ? (size_t)(d / 2) * ld : 0)) * sizeof(float);
        int tpc = 8;
        while (d * tpc > kWideBwdThreads) tpc >>= 1;
        static DeviceOnce wide_attr;                   // the opt-in is per device
        if (!wide_attr.done()) {
            cudaError_t e = cudaFuncSetAttribute(chol_wide_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e != cudaSuccess) { set_error("chol_bwd attr: %s", cudaGetErrorString(e)); return (int)e; }
            wide_attr.mark();
        }
        chol_wide_bwd_kernel<<<(unsigned)B, kWideBwdThreads, smem, (cudaStream_t)stream>>>(d, ld, tpc, vec, blocked, L, G, g_logdet, g_l1, penalty, g_G);
        CMF_CHECK_LAUNCH();
        return 0;
    }
    const int tps = pick_tps(d), spb = kCholThreads / tps;
    const size_t smem = (size_t)spb * 2 * d * (d + 1) * sizeof(float);
    static DeviceOnce attr_set;
    if (!attr_set.done()) {
        cudaError_t e = cudaFuncSetAttribute(chol_logdet_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) { set_error("chol_bwd attr: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set.mark();
    }
    chol_logdet_bwd_kernel<<<(unsigned)ceil_div(B, spb), kCholThreads, smem, (cudaStream_t)stream>>>(
        B, d, tps, L, G, g_logdet, g_l1, penalty, g_G);
    CMF_CHECK_LAUNCH();
    return 0;
}
