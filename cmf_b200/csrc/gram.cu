// Kernel (b), portable-precision path: per-sample metric G = V V^T (V = J^T, shape d x n) and its
// adjoint g_V = (g_G + g_G^T) V, as classic shared-memory tiled FP32 FMA GEMMs (64x64x16 tiles,
// 4x4 register micro-tiles).  Used for every d; the tcgen05 3xTF32 kernel (gram_tc.cu) takes over
// for the tensor-core friendly shapes.  Only the lower-triangular tiles of G are computed and the
// mirror is written from the same registers (SYRK).
#include "common.cuh"

namespace cmf {

constexpr int BT = 64;       // tile edge in M and N
constexpr int BK = 16;       // k-slab
constexpr int LDS_ = BT + 4; // padded smem row (keeps float4 alignment)
constexpr int kGemmThreads = 256;

// load a [64 rows x 16 k] slab of a row-major matrix (ld = row stride, k contiguous) transposed
// into s[k][row]; rows >= nrows / k >= nk read as zero.
__device__ __forceinline__ void load_slab_T(float (*s)[LDS_], const float* __restrict__ base, int ld,
                                            int row0, int nrows, int k0, int nk) {
    const int row = threadIdx.x >> 2;
    const int kq = (threadIdx.x & 3) << 2;
    const int r = row0 + row;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int k = k0 + kq + i;
        s[kq + i][row] = (r < nrows && k < nk) ? __ldg(base + (size_t)r * ld + k) : 0.f;
    }
}

__device__ __forceinline__ void fma_tile(float (&acc)[4][4], const float (*As)[LDS_], const float (*Bs)[LDS_],
                                         int ty, int tx) {
#pragma unroll
    for (int k = 0; k < BK; ++k) {
        const float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
        const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
        const float av[4] = {a.x, a.y, a.z, a.w};
        const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
}

// grid.x = B * T(T+1)/2 (sample-major; lower-triangular tile pairs)
__global__ void __launch_bounds__(kGemmThreads)
gram_fwd_kernel(int d, int n, int ntiles, const float* __restrict__ V, float* __restrict__ G, int accumulate) {
    __shared__ __align__(16) float As[BK][LDS_];
    __shared__ __align__(16) float Bs[BK][LDS_];
    const int b = blockIdx.x / ntiles;
    int tr = 0, rem = blockIdx.x - b * ntiles;
    while (rem > tr) { rem -= tr + 1; ++tr; }   // tile id -> (tr, tc), tc <= tr
    const int tc = rem;
    const float* Vb = V + (size_t)b * d * n;
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < n; k0 += BK) {
        load_slab_T(As, Vb, n, tr * BT, d, k0, n);
        load_slab_T(Bs, Vb, n, tc * BT, d, k0, n);
        __syncthreads();
        fma_tile(acc, As, Bs, ty, tx);
        __syncthreads();
    }
    float* Gb = G + (size_t)b * d * d;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = tr * BT + ty * 4 + i;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = tc * BT + tx * 4 + j;
            if (r < d && c < d) {
                const float v = acc[i][j];
                if (accumulate) {
                    Gb[(size_t)r * d + c] += v;
                    if (tr != tc) Gb[(size_t)c * d + r] += v;
                } else {
                    Gb[(size_t)r * d + c] = v;
                    if (tr != tc) Gb[(size_t)c * d + r] = v;
                }
            }
        }
    }
}

// g_V[b] = (g_G[b] + g_G[b]^T) V[b].  grid = (B * ceil(n/64), ceil(d/64))
__global__ void __launch_bounds__(kGemmThreads)
gram_bwd_kernel(int d, int n, int nx, const float* __restrict__ V, const float* __restrict__ gG,
                float* __restrict__ gV) {
    __shared__ __align__(16) float As[BK][LDS_];
    __shared__ __align__(16) float Bs[BK][LDS_];
    const int b = blockIdx.x / nx;
    const int r0 = blockIdx.y * BT, n0 = (blockIdx.x - b * nx) * BT;
    const float* Vb = V + (size_t)b * d * n;
    const float* Sb = gG + (size_t)b * d * d;
    const int ty = threadIdx.x >> 4, tx = threadIdx.x & 15;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < d; k0 += BK) {
        {   // As[k][m] = S[m][k] + S[k][m]
            const int row = threadIdx.x >> 2, kq = (threadIdx.x & 3) << 2;
            const int m = r0 + row;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int k = k0 + kq + i;
                As[kq + i][row] = (m < d && k < d) ? __ldg(Sb + (size_t)m * d + k) + __ldg(Sb + (size_t)k * d + m) : 0.f;
            }
        }
        {   // Bs[k][nn] = V[k][n0 + nn]
            const int kk = threadIdx.x >> 4, nq = (threadIdx.x & 15) << 2;
            const int k = k0 + kk;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int c = n0 + nq + i;
                Bs[kk][nq + i] = (k < d && c < n) ? __ldg(Vb + (size_t)k * n + c) : 0.f;
            }
        }
        __syncthreads();
        fma_tile(acc, As, Bs, ty, tx);
        __syncthreads();
    }
    float* Ob = gV + (size_t)b * d * n;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = r0 + ty * 4 + i;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = n0 + tx * 4 + j;
            if (r < d && c < n) Ob[(size_t)r * n + c] = acc[i][j];
        }
    }
}


// ---------------------------------------------------------------------------------------------
// Adjoint for d <= 64 (the MNIST-shape latent): one CTA per (sample, group of 128-column chunks).
// S + S^T (64 x 64, zero padded) is staged ONCE per CTA in shared memory; V chunks of 64 x 128
// arrive by 16-byte cp.async into a double buffer while the previous chunk is multiplied.  A warp
// owns 16 output columns and all 64 rows (8 x 4 register tile per thread: two broadcast-free 128-bit
// loads of S and one of V per k, 32 FMAs), so the ragged last chunk (n = 392 = 3 x 128 + 8) costs
// one warp, not a tile.  The generic kernel above re-reads S (transposed, uncoalesced) for every
// 64 x 64 output tile and is bound by that: 101 us at B = 512, n = 392 against an FMA floor of 25.
// ---------------------------------------------------------------------------------------------
constexpr int kSmallD = 64;
constexpr int kChunkN = 128;
constexpr int kALd = kSmallD + 4;        // 68: rows 16-byte aligned
constexpr int kBLd = kChunkN + 4;        // 132
constexpr size_t kSmallSmem = (size_t)(kSmallD * kALd + 2 * kSmallD * kBLd) * sizeof(float);

__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc)
                 : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kN>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(kN) : "memory"); }

// grid = (B, splits); CTA (b, s) handles chunks s, s + splits, ... of sample b
__global__ void __launch_bounds__(256, 2)
gram_bwd_small_kernel(int d, int n, const float* __restrict__ V, const float* __restrict__ gG, float* __restrict__ gV) {
    extern __shared__ __align__(16) float sm[];
    float* As = sm;                                   // [64][68]   As[k][m] = S[k][m] + S[m][k]
    float* Bs = sm + kSmallD * kALd;                  // [2][64][132]
    const int b = blockIdx.x, tid = threadIdx.x;
    const float* Vb = V + (size_t)b * d * n;
    const float* Sb = gG + (size_t)b * d * d;
    float* Ob = gV + (size_t)b * d * n;
    const int nchunks = (n + kChunkN - 1) / kChunkN;

    auto load_chunk = [&](int c, int buf) {           // V[0..d) x [c*128, c*128+128) -> Bs[buf]
        float* dst = Bs + (size_t)buf * kSmallD * kBLd;
        const int c0 = c * kChunkN;
        for (int e = tid; e < d * (kChunkN / 4); e += 256) {
            const int k = e >> 5, q = (e & 31) << 2;
            if (c0 + q < n) cp_async16(dst + k * kBLd + q, Vb + (size_t)k * n + c0 + q);
        }
        cp_async_commit();
    };

    int c = blockIdx.y;
    if (c < nchunks) load_chunk(c, 0);
    // stage S raw into the second V buffer, then symmetrise into As (zero padded to 64 x 64)
    {
        float* T = Bs + (size_t)kSmallD * kBLd;       // [64][65]
        float sv[kSmallD * kSmallD / 256];            // all 16 loads in flight before the first store
#pragma unroll
        for (int u = 0; u < kSmallD * kSmallD / 256; ++u) {
            const int e = tid + 256 * u, r = e >> 6, q = e & 63;
            sv[u] = (r < d && q < d) ? __ldg(Sb + (size_t)r * d + q) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < kSmallD * kSmallD / 256; ++u) {
            const int e = tid + 256 * u;
            T[(e >> 6) * 65 + (e & 63)] = sv[u];
        }
        __syncthreads();
        for (int e = tid; e < kSmallD * kSmallD; e += 256) {
            const int k = e >> 6, m = e & 63;
            As[k * kALd + m] = T[k * 65 + m] + T[m * 65 + k];
        }
        __syncthreads();
    }
    const int warp = tid >> 5, lane = tid & 31;
    const int mg = lane >> 2, cg = lane & 3;          // rows {4 mg .. +3} and {32 + 4 mg .. +3}; columns 16 warp + 4 cg .. +3
    int buf = 0;
    for (; c < nchunks; c += gridDim.y, buf ^= 1) {
        const int cn = c + gridDim.y;
        if (cn < nchunks) { load_chunk(cn, buf ^ 1); cp_async_wait<1>(); } else cp_async_wait<0>();
        __syncthreads();
        const int col0 = c * kChunkN + 16 * warp;
        if (col0 < n) {                               // warp-uniform: the ragged chunk runs on its live warps only
            const float* Bc = Bs + (size_t)buf * kSmallD * kBLd + 16 * warp + 4 * cg;
            float acc[8][4];
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 4
            for (int k = 0; k < d; ++k) {
                const float4 a0 = *reinterpret_cast<const float4*>(As + k * kALd + 4 * mg);
                const float4 a1 = *reinterpret_cast<const float4*>(As + k * kALd + 32 + 4 * mg);
                const float4 v = *reinterpret_cast<const float4*>(Bc + k * kBLd);
                const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
                const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int i = 0; i < 8; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], vv[j], acc[i][j]);
            }
            const int col = col0 + 4 * cg;
            if (col < n) {
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int m = (i < 4 ? 4 * mg + i : 32 + 4 * mg + (i - 4));
                    if (m < d) stcs4(Ob + (size_t)m * n + col, make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]));
                }
            }
        }
        __syncthreads();                              // everyone is done with Bs[buf] before it is refilled
    }
}

// tensor-core path (gram_tc.cu)
bool gram_tc_usable(int B, int d, int n, const float* V, const float* G);
int gram_tc_fwd_launch(int B, int d, int n, const float* V, int n2, const float* V2, float* G, int accumulate,
                       cudaStream_t stream);
bool gram_tc_bwd_usable(int B, int d, int n, const float* V, const float* gG, const float* gV);
int gram_tc_bwd_launch(int B, int d, int n, const float* V, const float* gG, float* gV, cudaStream_t stream);

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_gram_fwd(int B, int d, int n, const float* V, float* G, int accumulate, cmf_stream_t stream) {
    return cmf_gram_fwd_path(B, d, n, V, G, accumulate, CMF_GRAM_AUTO, stream);
}

extern "C" int cmf_gram_fwd_path(int B, int d, int n, const float* V, float* G, int accumulate, int path,
                                 cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && d >= 0 && n >= 0, "gram_fwd: negative size");
    CMF_REQUIRE(path >= CMF_GRAM_AUTO && path <= CMF_GRAM_TC, "gram_fwd: bad path %d", path);
    if (B == 0 || d == 0) return 0;
    CMF_REQUIRE(G && (n == 0 || V), "gram_fwd: null pointer");
    if (path != CMF_GRAM_FMA && n > 0) {
        const bool ok = gram_tc_usable(B, d, n, V, G);
        CMF_REQUIRE(ok || path == CMF_GRAM_AUTO,
                    "gram_fwd: tensor-core path needs d >= 16, n >= 32, n %% 4 == 0 and 16-byte aligned V (d=%d n=%d)", d, n);
        if (ok) return gram_tc_fwd_launch(B, d, n, V, 0, nullptr, G, accumulate, (cudaStream_t)stream);
    }
    if (n == 0) {
        if (!accumulate) {
            cudaError_t e = cudaMemsetAsync(G, 0, (size_t)B * d * d * sizeof(float), (cudaStream_t)stream);
            if (e != cudaSuccess) { set_error("gram_fwd memset: %s", cudaGetErrorString(e)); return (int)e; }
        }
        return 0;
    }
    const int T = (int)ceil_div(d, BT);
    const int ntiles = T * (T + 1) / 2;
    CMF_REQUIRE((long long)B * ntiles < (1LL << 31), "gram_fwd: grid too large");
    gram_fwd_kernel<<<(unsigned)(B * ntiles), kGemmThreads, 0, (cudaStream_t)stream>>>(d, n, ntiles, V, G, accumulate);
    CMF_CHECK_LAUNCH();
    return 0;
}

extern "C" int cmf_gram_fwd2(int B, int d, int n1, const float* V1, int n2, const float* V2, float* G, int accumulate,
                             cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && d >= 0 && n1 >= 0 && n2 >= 0, "gram_fwd2: negative size");
    if (B == 0 || d == 0) return 0;
    CMF_REQUIRE(G && (n1 == 0 || V1) && (n2 == 0 || V2), "gram_fwd2: null pointer");
    if (n1 > 0 && n2 > 0 && gram_tc_usable(B, d, n1, V1, G) && gram_tc_usable(B, d, n2, V2, G))
        return gram_tc_fwd_launch(B, d, n1, V1, n2, V2, G, accumulate, (cudaStream_t)stream);
    int rc = cmf_gram_fwd_path(B, d, n1, V1, G, accumulate, CMF_GRAM_AUTO, stream);
    if (rc != 0) return rc;
    return cmf_gram_fwd_path(B, d, n2, V2, G, 1, CMF_GRAM_AUTO, stream);
}

extern "C" int cmf_gram_bwd(int B, int d, int n, const float* V, const float* g_G, float* g_V, cmf_stream_t stream) {
    return cmf_gram_bwd_path(B, d, n, V, g_G, g_V, CMF_GRAM_AUTO, stream);
}

extern "C" int cmf_gram_bwd_path(int B, int d, int n, const float* V, const float* g_G, float* g_V, int path,
                                 cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && d >= 0 && n >= 0, "gram_bwd: negative size");
    CMF_REQUIRE(path >= CMF_GRAM_AUTO && path <= CMF_GRAM_TC, "gram_bwd: bad path %d", path);
    if (B == 0 || d == 0 || n == 0) return 0;
    CMF_REQUIRE(V && g_G && g_V, "gram_bwd: null pointer");
    if (path != CMF_GRAM_FMA) {
        const bool ok = gram_tc_bwd_usable(B, d, n, V, g_G, g_V);
        CMF_REQUIRE(ok || path == CMF_GRAM_AUTO,
                    "gram_bwd: tensor-core path needs d >= 16, d %% 4 == 0, n >= 32, n %% 4 == 0 and 16-byte aligned V, g_G (d=%d n=%d)", d, n);
        // measured (scripts/gram_probe.py): the persistent tensor-core kernel wins from d = 32 up, the FP32 FMA kernel at d = 16
        if (ok && (path == CMF_GRAM_TC || d >= 32)) return gram_tc_bwd_launch(B, d, n, V, g_G, g_V, (cudaStream_t)stream);
    }
    if (d <= kSmallD && n % 4 == 0 && aligned16(V) && aligned16(g_V) && B <= 65535 * 32) {
        static DeviceOnce attr_set;
        if (!attr_set.done()) {
            cudaError_t e = cudaFuncSetAttribute(gram_bwd_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmallSmem);
            if (e != cudaSuccess) { set_error("gram_bwd attr: %s", cudaGetErrorString(e)); return (int)e; }
            attr_set.mark();
        }
        // split a sample's chunks over several CTAs while the grid is below ~2 CTAs per SM
        const int nchunks = (int)ceil_div(n, kChunkN);
        int splits = 1;
        while (splits < nchunks && (long long)B * splits < 2 * kNumSMs) ++splits;
        dim3 grid((unsigned)B, (unsigned)splits);
        gram_bwd_small_kernel<<<grid, 256, kSmallSmem, (cudaStream_t)stream>>>(d, n, V, g_G, g_V);
        CMF_CHECK_LAUNCH();
        return 0;
    }
    const int nx = (int)ceil_div(n, BT);
    CMF_REQUIRE((long long)B * nx < (1LL << 31), "gram_bwd: grid too large");
    dim3 grid((unsigned)(B * nx), (unsigned)ceil_div(d, BT));
    gram_bwd_kernel<<<grid, kGemmThreads, 0, (cudaStream_t)stream>>>(d, n, nx, V, g_G, g_V);
    CMF_CHECK_LAUNCH();
    return 0;
}
