// Kernel (b), tensor-core path: per-sample metric G = V V^T (V = J^T, (d x n) row-major, K-major for
// both operands) on tcgen05 with TMEM accumulators, FP32-accurate by the 3xTF32 split
//     V = hi + lo   (hi = fp32 with the 13 low mantissa bits cleared, lo = V - hi exactly)
//     G ~= hi hi^T + hi lo^T + lo hi^T            (the dropped lo lo^T term is O(2^-22) relative)
// replacing torch.stack + torch.bmm of non_square.py:307-308 (cuBLAS batched SGEMM).
//
// Tiling.  The tangents of the whole batch are one (B*d) x n matrix.  A CTA owns one 128 x 128
// accumulator tile in TMEM:
//   d <= 128 : a "panel" of S = floor(128/d) whole samples (S*d rows) is both operands; the S Gram
//              matrices are the diagonal d x d blocks of the tile (the cross-sample blocks are
//              computed and dropped: an M=64 MMA costs the same tensor cycles as M=128).
//   d  > 128 : tiles (tr, tc <= tr) of one sample, SYRK style; the mirror is written from the same
//              accumulator, diagonal tiles write their lower triangle twice, so G is bitwise symmetric.
// Pipeline (192 threads): warp 0 = TMA producer (one lane; 128-byte-swizzled boxes of 32 k-columns),
// warps 2..5 = split the landed tile in shared memory (hi in place, lo beside it) and later the
// epilogue (tcgen05.ld -> global), warp 1 = TMEM allocation + MMA issue (one lane).
// HBM traffic: V read once (4 d n bytes / sample) + G written; 3 x 2 d^2 n tensor FLOP / sample.
#include "common.cuh"
#include "tc_common.cuh"

namespace cmf {

using namespace tc;

constexpr int kTile = 128;            // accumulator tile edge (UMMA M = N = 128)
constexpr int kKc = 32;               // k-columns per stage: 32 fp32 = one 128-byte swizzle row
constexpr int kTileBytes = kTile * kKc * 4;   // 16 KB
constexpr int kMaxStages = 4;
constexpr int kTmemCols = 512;        // [0,128) cross terms, [128,256) / [256,384) hi*hi ping-pong
constexpr int kSplitWarps = 4, kDrainWarps = 8;
constexpr int kGramTcThreads = 32 * (2 + kSplitWarps + kDrainWarps);   // 448

struct GramTcParams {
    int B, d, n;
    int S;          // samples per panel (d <= 128) or 0 (d > 128)
    int T;          // 128-row tiles per sample (d > 128)
    int ntiles;     // T (T+1) / 2
    int nstages;
    int stage_bytes;
    int box_rows;
    int accumulate;
    float* G;
};

__device__ __forceinline__ void g_store(float* p, float v, int accumulate) {
    *p = accumulate ? (*p + v) : v;
}

// Accumulation precision.  The tensor core adds into its FP32 accumulator with truncation, so a
// long K loop into one accumulator picks up a bias of ~0.3 ulp per MMA (measured 1.3e-5 relative
// at K = 784 -- cond(G) times that is far outside the 1e-4 parity contract).  Therefore:
//   * the large term hi*hi^T is accumulated on the tensor core only over ONE stage (4 MMAs, K=32)
//     into a ping-pong TMEM buffer, then drained and added to FP32 registers with round-to-nearest
//     (8 drain warps, 64 accumulators per thread) while the next stage's MMAs run;
//   * the cross terms hi*lo^T + lo*hi^T (2^-11 of the magnitude) stay in one TMEM accumulator for
//     the whole K loop -- their truncation error is 2^-11 of an ulp of G per step.
__global__ void __launch_bounds__(kGramTcThreads, 1)
gram_tc_fwd_kernel(const __grid_constant__ CUtensorMap tmap, const GramTcParams P) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kMaxStages], split_bar[kMaxStages], empty_bar[kMaxStages];
    __shared__ __align__(8) uint64_t hh_full[2], hh_free[2], done_bar;
    __shared__ uint32_t tmem_base_smem;

    // 1024-byte alignment for the 128-byte swizzle atoms
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // ---- work item
    int rowA, rowB, b0, tr = 0, tc = 0;
    bool same;
    if (P.S > 0) {
        b0 = blockIdx.x * P.S;
        rowA = rowB = b0 * P.d;
        same = true;
    } else {
        b0 = blockIdx.x / P.ntiles;
        int rem = blockIdx.x - b0 * P.ntiles;
        while (rem > tr) { rem -= tr + 1; ++tr; }
        tc = rem;
        rowA = b0 * P.d + tr * kTile;
        rowB = b0 * P.d + tc * kTile;
        same = (tr == tc);
    }
    const int nk = (P.n + kKc - 1) / kKc;
    const int nst = P.nstages;

    if (threadIdx.x == 0) {
        for (int s = 0; s < nst; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&split_bar[s], 32 * kSplitWarps);
            mbar_init(&empty_bar[s], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&hh_full[i], 1);
            mbar_init(&hh_free[i], 32 * kDrainWarps);
        }
        mbar_init(&done_bar, 1);
        mbar_fence_init();
        prefetch_tmap(&tmap);
    }
    if (warp == 1) tmem_alloc<kTmemCols>(&tmem_base_smem);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    if (warp == 0) {
        // ===== TMA producer
        if (lane == 0) {
            const uint32_t tx = (uint32_t)P.box_rows * kKc * 4 * (same ? 1 : 2);
            for (int kc = 0; kc < nk; ++kc) {
                const int s = kc % nst;
                const uint32_t ph = (uint32_t)(kc / nst) & 1u;
                mbar_wait(&empty_bar[s], ph ^ 1u);
                uint8_t* st = smem + (size_t)s * P.stage_bytes;
                mbar_arrive_expect_tx(&full_bar[s], tx);
                tma_load_2d(st, &tmap, &full_bar[s], kc * kKc, rowA);
                if (!same) tma_load_2d(st + 2 * kTileBytes, &tmap, &full_bar[s], kc * kKc, rowB);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_tf32(kTile, kTile);
            for (int kc = 0; kc < nk; ++kc) {
                const int s = kc % nst;
                const uint32_t ph = (uint32_t)(kc / nst) & 1u;
                const int pp = kc & 1;
                mbar_wait(&split_bar[s], ph);
                mbar_wait(&hh_free[pp], ((uint32_t)(kc >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t a_hi = smem_u32(smem + (size_t)s * P.stage_bytes);
                const uint32_t a_lo = a_hi + kTileBytes;
                const uint32_t b_hi = same ? a_hi : a_hi + 2 * kTileBytes;
                const uint32_t b_lo = b_hi + kTileBytes;
                const uint32_t t_hh = tmem_base + 128u + 128u * (uint32_t)pp;
#pragma unroll
                for (int k = 0; k < kKc / 8; ++k)            // UMMA K = 8 for tf32 (32 bytes)
                    umma_tf32(t_hh, umma_desc_sw128(a_hi + k * 32), umma_desc_sw128(b_hi + k * 32), idesc, k != 0);
                umma_commit(&hh_full[pp]);                   // drain warps may read hh[pp]
#pragma unroll
                for (int k = 0; k < kKc / 8; ++k) {
                    umma_tf32(tmem_base, umma_desc_sw128(a_lo + k * 32), umma_desc_sw128(b_hi + k * 32), idesc,
                              (kc | k) != 0);
                    umma_tf32(tmem_base, umma_desc_sw128(a_hi + k * 32), umma_desc_sw128(b_lo + k * 32), idesc, 1);
                }
                umma_commit(&empty_bar[s]);                  // smem slot free once these MMAs retire
            }
            umma_commit(&done_bar);                          // cross accumulator complete
        }
    } else if (warp < 2 + kSplitWarps) {
        // ===== splitter warps: hi in place, lo beside it
        const int t = threadIdx.x - 64;                      // 0..127
        const int nvec = (same ? 1 : 2) * (kTileBytes / 16); // float4 per stage of raw data
        for (int kc = 0; kc < nk; ++kc) {
            const int s = kc % nst;
            const uint32_t ph = (uint32_t)(kc / nst) & 1u;
            mbar_wait(&full_bar[s], ph);
            uint8_t* st = smem + (size_t)s * P.stage_bytes;
            for (int i = t; i < nvec; i += 32 * kSplitWarps) {
                // tile 0: raw at [0,16K) -> lo at [16K,32K); tile 1: raw at [32K,48K) -> lo at [48K,64K)
                const int tile = i / (kTileBytes / 16), w = i - tile * (kTileBytes / 16);
                float4* raw = reinterpret_cast<float4*>(st + (size_t)tile * 2 * kTileBytes) + w;
                float4* lo = raw + kTileBytes / 16;
                const float4 v = *raw;
                float4 h, l;
                split_tf32(v.x, h.x, l.x); split_tf32(v.y, h.y, l.y);
                split_tf32(v.z, h.z, l.z); split_tf32(v.w, h.w, l.w);
                *raw = h;
                *lo = l;
            }
            fence_proxy_async_smem();
            mbar_arrive(&split_bar[s]);
        }
    } else {
        // ===== drain warps: FP32 round-to-nearest accumulation of hi*hi^T, then the epilogue
        const int dw = warp - (2 + kSplitWarps);             // 0..7
        const int q = warp & 3;                              // TMEM lane quarter this warp may access
        const int half = dw >> 2;                            // column half [64*half, 64*half + 64)
        const int i = q * 32 + lane;                         // tile row
        const int d = P.d;
        // does this warp's (32 rows x 64 columns) window hold anything that is written out?
        bool needed = true;
        if (P.S > 0) {
            const int r0 = q * 32, r1 = q * 32 + 31;
            const int c_lo = (r0 / d) * d, c_hi = min((r1 / d) * d + d, P.S * d);
            needed = (r0 < P.S * d) && (half * 64 < c_hi) && (half * 64 + 64 > c_lo);
        }
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        float acc[2][32];
#pragma unroll
        for (int ch = 0; ch < 2; ++ch)
#pragma unroll
            for (int u = 0; u < 32; ++u) acc[ch][u] = 0.f;
        for (int kc = 0; kc < nk; ++kc) {
            const int pp = kc & 1;
            mbar_wait(&hh_full[pp], (uint32_t)(kc >> 1) & 1u);
            tc_fence_after();
            if (needed) {
                const uint32_t t_hh = tmem_base + 128u + 128u * (uint32_t)pp + (uint32_t)(half * 64);
#pragma unroll
                for (int ch = 0; ch < 2; ++ch) {
                    float v[32];
                    tmem_ld_32x32(t_hh + lane_addr + (uint32_t)(ch * 32), v);
#pragma unroll
                    for (int u = 0; u < 32; ++u) acc[ch][u] += v[u];
                }
            }
            tc_fence_before();
            mbar_arrive(&hh_free[pp]);
        }
        // ---- epilogue: + cross terms, TMEM/registers -> global
        mbar_wait(&done_bar, 0);
        tc_fence_after();
        if (needed) {
            float* G = P.G;
            const int accm = P.accumulate;
#pragma unroll
            for (int ch = 0; ch < 2; ++ch) {
                const int c0 = half * 64 + ch * 32;
                float v[32];
                tmem_ld_32x32(tmem_base + lane_addr + (uint32_t)c0, v);
                if (P.S > 0) {
                    const int s = i / d, li = i - s * d, b = b0 + s;
                    if (s < P.S && b < P.B) {
                        float* Gb = G + (size_t)b * d * d;
#pragma unroll
                        for (int u = 0; u < 32; ++u) {
                            const int j = c0 + u - s * d;
                            if (j >= 0 && j <= li) {
                                const float g = acc[ch][u] + v[u];
                                g_store(Gb + (size_t)li * d + j, g, accm);
                                if (j < li) g_store(Gb + (size_t)j * d + li, g, accm);
                            }
                        }
                    }
                } else {
                    const int gi = tr * kTile + i;
                    if (gi < d) {
                        float* Gb = G + (size_t)b0 * d * d;
#pragma unroll
                        for (int u = 0; u < 32; ++u) {
                            const int gj = tc * kTile + c0 + u;
                            if (gj < d && (tr != tc || gj <= gi)) {
                                const float g = acc[ch][u] + v[u];
                                g_store(Gb + (size_t)gi * d + gj, g, accm);
                                if (gj != gi) g_store(Gb + (size_t)gj * d + gi, g, accm);
                            }
                        }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kTmemCols>(tmem_base);
    }
}

// ------------------------------------------------------------------------------------------------
// host
// ------------------------------------------------------------------------------------------------
cmf_encode_tiled_fn get_encode_tiled() {
    static cmf_encode_tiled_fn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<cmf_encode_tiled_fn>(p);
    }
    return fn;
}

int make_tmap_2d_f32(CUtensorMap* out, const float* base, uint64_t rows, uint64_t cols, uint64_t ld,
                     uint32_t box_rows) {
    cmf_encode_tiled_fn enc = get_encode_tiled();
    if (!enc) { set_error("cuTensorMapEncodeTiled unavailable"); return -2; }
    const cuuint64_t gdim[2] = {cols, rows};
    const cuuint64_t gstride[1] = {ld * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)kKc, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return -3; }
    return 0;
}

bool gram_tc_usable(int B, int d, int n, const float* V, const float* G) {
    return B > 0 && d >= 16 && n >= kKc && (n & 3) == 0 && aligned16(V) && G != nullptr &&
           (long long)B * d < (1LL << 31);
}

int gram_tc_fwd_launch(int B, int d, int n, const float* V, float* G, int accumulate, cudaStream_t stream) {
    GramTcParams P{};
    P.B = B; P.d = d; P.n = n; P.G = G; P.accumulate = accumulate;
    long long grid;
    if (d <= kTile) {
        P.S = kTile / d; P.T = 1; P.ntiles = 1;
        P.box_rows = P.S * d;
        P.nstages = 4; P.stage_bytes = 2 * kTileBytes;
        grid = ceil_div(B, P.S);
    } else {
        P.S = 0; P.T = (int)ceil_div(d, kTile); P.ntiles = P.T * (P.T + 1) / 2;
        P.box_rows = kTile;
        P.nstages = 3; P.stage_bytes = 4 * kTileBytes;
        grid = (long long)B * P.ntiles;
    }
    CMF_REQUIRE(grid < (1LL << 31), "gram_tc: grid too large");
    CUtensorMap tmap;
    int rc = make_tmap_2d_f32(&tmap, V, (uint64_t)B * d, (uint64_t)n, (uint64_t)n, (uint32_t)P.box_rows);
    if (rc != 0) return rc;
    const size_t smem = (size_t)P.nstages * P.stage_bytes + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(gram_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) { set_error("gram_tc attr: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set = true;
    }
    gram_tc_fwd_kernel<<<(unsigned)grid, kGramTcThreads, smem, stream>>>(tmap, P);
    CMF_CHECK_LAUNCH();
    return 0;
}

}  // namespace cmf
