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
// Pipeline (448 threads): warp 0 = TMA producer (one lane; 128-byte-swizzled boxes of 32 k-columns),
// warp 1 = TMEM allocation + MMA issue (one lane), warps 2..5 = split the landed tile in shared
// memory (hi in place, lo beside it), warps 6..13 = drain (see "Accumulation precision") + epilogue.
// HBM traffic: V read once (4 d n bytes / sample) + G written; 3 x 2 d^2 n tensor FLOP / sample.
#include "common.cuh"
#include "tc_common.cuh"

namespace cmf {

using namespace tc;

constexpr int kTile = 128;            // accumulator tile edge (UMMA M = N = 128)
constexpr int kKc = 32;               // k-columns per stage: 32 fp32 = one 128-byte swizzle row
constexpr int kTileBytes = kTile * kKc * 4;   // 16 KB
constexpr int kMaxStages = 10;
constexpr int kTmemCols = 512;        // [0,128) cross terms, [128,256) / [256,384) hi*hi ping-pong
constexpr int kSplitWarps = 4, kDrainWarps = 8;
constexpr int kDrainEvery = 4;        // stages (of K = 32) accumulated on the tensor core per drain
constexpr int kGramTcThreads = 32 * (2 + kSplitWarps + kDrainWarps);   // 448
constexpr int kGsLd = kTile + 1;      // padded row of the epilogue staging tile (bank-conflict free)

struct GramTcParams {
    int B, d, n;
    int n2;         // columns of the optional second block (0: none)
    int S;          // samples per panel (d <= 128) or 0 (d > 128)
    int T;          // 128-row tiles per sample (d > 128)
    int ntiles;     // T (T+1) / 2
    int nstages;    // raw ring depth
    int raw_bytes;  // bytes of one raw stage (one or two 16 KB tiles)
    int box_rows;
    int accumulate;
    float* G;
};

// Accumulation precision.  The tensor core adds into its FP32 accumulator with truncation, so a
// long K loop into one accumulator picks up a bias of ~0.3-0.5 ulp per MMA (measured 1.3e-5
// relative at K = 784 -- cond(G) times that is far outside the 1e-4 parity contract).  Therefore:
//   * the large term hi*hi^T is accumulated on the tensor core only over kDrainEvery stages
//     (16 MMAs, K = 128) into a ping-pong TMEM buffer, then drained and added to FP32 registers
//     with round-to-nearest (8 drain warps, 64 accumulators per thread) while the next group runs;
//   * the cross terms hi*lo^T + lo*hi^T (2^-11 of the magnitude) stay in one TMEM accumulator for
//     the whole K loop -- their truncation error is 2^-11 of an ulp of G per step.
// Shared memory.  The pipeline is latency-bound (TMA latency + split + MMA per stage), so the raw
// ring is deep (8 x 16 KB, or 4 x 32 KB when the two operands differ) and the lo halves live in a
// separate 2-deep ring that only spans split -> MMA.  After the K loop the ring is reused as a
// 128 x 129 staging tile so that G leaves the SM in coalesced rows (and mirrored rows).
__global__ void __launch_bounds__(kGramTcThreads, 1)
gram_tc_fwd_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap2,
                   const GramTcParams P) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kMaxStages], empty_bar[kMaxStages];
    __shared__ __align__(8) uint64_t split_bar[2], lo_free[2], hh_full[2], hh_free[2], done_bar;
    __shared__ uint32_t tmem_base_smem;

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
    // K loop over the columns of the first block, then (n2 > 0) of a second block of the same
    // samples: G = V1 V1^T + V2 V2^T in one accumulation, one launch (the two halves of J^T)
    const int nk1 = (P.n + kKc - 1) / kKc;
    const int nk = nk1 + (P.n2 + kKc - 1) / kKc;
    const int nst = P.nstages;
    uint8_t* lo_ring = smem + (size_t)nst * P.raw_bytes;      // 2 x raw_bytes

    if (threadIdx.x == 0) {
        for (int s = 0; s < nst; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&split_bar[i], 32 * kSplitWarps);
            mbar_init(&lo_free[i], 1);
            mbar_init(&hh_full[i], 1);
            mbar_init(&hh_free[i], 32 * kDrainWarps);
        }
        mbar_init(&done_bar, 1);
        mbar_fence_init();
        prefetch_tmap(&tmap);
        if (P.n2 > 0) prefetch_tmap(&tmap2);
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
                uint8_t* st = smem + (size_t)s * P.raw_bytes;
                mbar_arrive_expect_tx(&full_bar[s], tx);
                const CUtensorMap* tm = kc < nk1 ? &tmap : &tmap2;
                const int col = (kc < nk1 ? kc : kc - nk1) * kKc;
                tma_load_2d(st, tm, &full_bar[s], col, rowA);
                if (!same) tma_load_2d(st + kTileBytes, tm, &full_bar[s], col, rowB);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_tf32(kTile, kTile);
            for (int kc = 0; kc < nk; ++kc) {
                const int s = kc % nst, l = kc & 1;
                const int g = kc / kDrainEvery, r = kc - g * kDrainEvery, pp = g & 1;
                mbar_wait(&split_bar[l], (uint32_t)(kc >> 1) & 1u);
                if (r == 0) mbar_wait(&hh_free[pp], ((uint32_t)(g >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t a_hi = smem_u32(smem + (size_t)s * P.raw_bytes);
                const uint32_t a_lo = smem_u32(lo_ring + (size_t)l * P.raw_bytes);
                const uint32_t b_hi = same ? a_hi : a_hi + kTileBytes;
                const uint32_t b_lo = same ? a_lo : a_lo + kTileBytes;
                const uint32_t t_hh = tmem_base + 128u + 128u * (uint32_t)pp;
#pragma unroll
                for (int k = 0; k < kKc / 8; ++k)            // UMMA K = 8 for tf32 (32 bytes)
                    umma_tf32(t_hh, umma_desc_sw128(a_hi + k * 32), umma_desc_sw128(b_hi + k * 32), idesc, (r | k) != 0);
                if (r == kDrainEvery - 1 || kc == nk - 1) umma_commit(&hh_full[pp]);   // drain warps may read hh[pp]
#pragma unroll
                for (int k = 0; k < kKc / 8; ++k) {
                    umma_tf32(tmem_base, umma_desc_sw128(a_lo + k * 32), umma_desc_sw128(b_hi + k * 32), idesc,
                              (kc | k) != 0);
                    umma_tf32(tmem_base, umma_desc_sw128(a_hi + k * 32), umma_desc_sw128(b_lo + k * 32), idesc, 1);
                }
                umma_commit(&empty_bar[s]);                  // raw slot and lo slot free once these retire
                umma_commit(&lo_free[l]);
            }
            umma_commit(&done_bar);                          // cross accumulator complete
        }
    } else if (warp < 2 + kSplitWarps) {
        // ===== splitter warps: hi in place, lo into the 2-deep lo ring
        const int t = threadIdx.x - 64;                      // 0..127
        const int nvec = (same ? 1 : 2) * (kTileBytes / 16); // float4 per stage of raw data
        for (int kc = 0; kc < nk; ++kc) {
            const int s = kc % nst, l = kc & 1;
            mbar_wait(&full_bar[s], (uint32_t)(kc / nst) & 1u);
            mbar_wait(&lo_free[l], ((uint32_t)(kc >> 1) & 1u) ^ 1u);
            float4* raw = reinterpret_cast<float4*>(smem + (size_t)s * P.raw_bytes);
            float4* lo = reinterpret_cast<float4*>(lo_ring + (size_t)l * P.raw_bytes);
#pragma unroll 4
            for (int i = t; i < nvec; i += 32 * kSplitWarps) {
                const float4 v = raw[i];
                float4 h, w;
                split_tf32(v.x, h.x, w.x); split_tf32(v.y, h.y, w.y);
                split_tf32(v.z, h.z, w.z); split_tf32(v.w, h.w, w.w);
                raw[i] = h;
                lo[i] = w;
            }
            fence_proxy_async_smem();
            mbar_arrive(&split_bar[l]);
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
        const int ngroups = (nk + kDrainEvery - 1) / kDrainEvery;
        for (int g = 0; g < ngroups; ++g) {
            const int pp = g & 1;
            mbar_wait(&hh_full[pp], (uint32_t)(g >> 1) & 1u);
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
        // ---- epilogue: + cross terms -> staging tile in the (now idle) ring -> coalesced global
        mbar_wait(&done_bar, 0);
        tc_fence_after();
        float* Gs = reinterpret_cast<float*>(smem);          // [128][kGsLd]
        if (needed) {
#pragma unroll
            for (int ch = 0; ch < 2; ++ch) {
                const int c0 = half * 64 + ch * 32;
                float v[32];
                tmem_ld_32x32(tmem_base + lane_addr + (uint32_t)c0, v);
#pragma unroll
                for (int u = 0; u < 32; ++u) Gs[i * kGsLd + c0 + u] = acc[ch][u] + v[u];
            }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kDrainWarps) : "memory");
        const int t = threadIdx.x - 32 * (2 + kSplitWarps);  // 0..255
        float* G = P.G;
        const int accm = P.accumulate;
        if (P.S > 0) {
            const int dd = d * d;
            const int total = min(P.S, P.B - b0) * dd;
            float* Gb = G + (size_t)b0 * dd;                 // the panel's samples are contiguous in G
            for (int e = t; e < total; e += 32 * kDrainWarps) {
                const int s = e / dd, rem = e - s * dd;
                const int li = rem / d, j = rem - li * d;
                const int hi_ = max(li, j), lo_ = min(li, j); // lower triangle is the source of both
                const float g = Gs[(s * d + hi_) * kGsLd + s * d + lo_];
                Gb[e] = accm ? Gb[e] + g : g;
            }
        } else {
            float* Gb = G + (size_t)b0 * d * d;
            for (int e = t; e < kTile * kTile; e += 32 * kDrainWarps) {
                const int r = e >> 7, c = e & 127;
                const int gi = tr * kTile + r, gj = tc * kTile + c;
                if (gi < d && gj < d) {
                    const float g = (tr != tc || c <= r) ? Gs[r * kGsLd + c] : Gs[c * kGsLd + r];
                    float* p = Gb + (size_t)gi * d + gj;
                    *p = accm ? *p + g : g;
                }
            }
            if (tr != tc) {                                  // mirror tile, rows of G^T coalesced
                for (int e = t; e < kTile * kTile; e += 32 * kDrainWarps) {
                    const int c = e >> 7, r = e & 127;
                    const int gi = tr * kTile + r, gj = tc * kTile + c;
                    if (gi < d && gj < d) {
                        const float g = Gs[r * kGsLd + c];
                        float* p = Gb + (size_t)gj * d + gi;
                        *p = accm ? *p + g : g;
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
                     uint32_t box_rows, int swizzle_atom32) {
    cmf_encode_tiled_fn enc = get_encode_tiled();
    if (!enc) { set_error("cuTensorMapEncodeTiled unavailable"); return -2; }
    const cuuint64_t gdim[2] = {cols, rows};
    const cuuint64_t gstride[1] = {ld * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)kKc, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE,
                     swizzle_atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return -3; }
    return 0;
}

// rank-2 fp32 tensor, box = box_rows x box_cols, no swizzle
int make_tmap_2d_plain_f32(CUtensorMap* out, const float* base, uint64_t rows, uint64_t cols, uint64_t ld,
                           uint32_t box_cols, uint32_t box_rows) {
    cmf_encode_tiled_fn enc = get_encode_tiled();
    if (!enc) { set_error("cuTensorMapEncodeTiled unavailable"); return -2; }
    const cuuint64_t gdim[2] = {cols, rows};
    const cuuint64_t gstride[1] = {ld * sizeof(float)};
    const cuuint32_t box[2] = {box_cols, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled (plain) failed (%d)", (int)r); return -3; }
    return 0;
}

bool gram_tc_usable(int B, int d, int n, const float* V, const float* G) {
    return B > 0 && d >= 16 && n >= kKc && (n & 3) == 0 && aligned16(V) && G != nullptr &&
           (long long)B * d < (1LL << 31);
}

int gram_tc_fwd_launch(int B, int d, int n, const float* V, int n2, const float* V2, float* G, int accumulate,
                       cudaStream_t stream) {
    GramTcParams P{};
    P.B = B; P.d = d; P.n = n; P.n2 = n2; P.G = G; P.accumulate = accumulate;
    long long grid;
    if (d <= kTile) {
        P.S = kTile / d; P.T = 1; P.ntiles = 1;
        P.box_rows = P.S * d;
        P.nstages = 8; P.raw_bytes = kTileBytes;
        grid = ceil_div(B, P.S);
    } else {
        P.S = 0; P.T = (int)ceil_div(d, kTile); P.ntiles = P.T * (P.T + 1) / 2;
        P.box_rows = kTile;
        P.nstages = 4; P.raw_bytes = 2 * kTileBytes;
        grid = (long long)B * P.ntiles;
    }
    CMF_REQUIRE(grid < (1LL << 31), "gram_tc: grid too large");
    CUtensorMap tmap;
    int rc = make_tmap_2d_f32(&tmap, V, (uint64_t)B * d, (uint64_t)n, (uint64_t)n, (uint32_t)P.box_rows, 0);
    if (rc != 0) return rc;
    CUtensorMap tmap2 = tmap;
    if (n2 > 0) {
        rc = make_tmap_2d_f32(&tmap2, V2, (uint64_t)B * d, (uint64_t)n2, (uint64_t)n2, (uint32_t)P.box_rows, 0);
        if (rc != 0) return rc;
    }
    const size_t smem = (size_t)(P.nstages + 2) * P.raw_bytes + 1024;     // raw ring + 2-deep lo ring
    static DeviceOnce attr_set;
    if (!attr_set.done()) {
        cudaError_t e = cudaFuncSetAttribute(gram_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) { set_error("gram_tc attr: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set.mark();
    }
    gram_tc_fwd_kernel<<<(unsigned)grid, kGramTcThreads, smem, stream>>>(tmap, tmap2, P);
    CMF_CHECK_LAUNCH();
    return 0;
}

}  // namespace cmf

// ================================================================================================
// Adjoint on the tensor core:  g_V[b] = (g_G[b] + g_G[b]^T) V[b]      (d x n), written transposed
// as the MMA  D[n, j] = sum_k V[k, n] * S[k, j],  S = g_G + g_G^T:
//   A = V tile, M = 128 n-positions, MN-major (n is contiguous in memory) -- TMA boxes of
//       32 n x 32 k land as 128-byte rows of k in the 32-byte-atom 128B swizzle
//       (CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B / UMMA SWIZZLE_128B_BASE32B), the only layout
//       tcgen05 accepts for an MN-major tf32 operand (the plain 128B swizzle yields zeros);
//   B = S tile, N = 16..128 output rows j, K-major (S is symmetric), built by the staging warps
//       from g_G (transpose-add + hi/lo split) directly in the swizzled layout.
// M = 128 is always full (no padding waste at d = 64).  K = d <= a few hundred MMAs, so the three
// 3xTF32 terms share one TMEM accumulator (truncation bias <~ 0.3 ulp per MMA, no drain needed).
// The epilogue reads a TMEM lane (= n) per thread, so every store of 32 lanes is one 128-byte line.
// ================================================================================================
namespace cmf {

constexpr int kBwdThreads = 192;      // warp 0 TMA, warp 1 MMA, warps 2..5 stage/split + epilogue
constexpr int kBwdTmemCols = 128;

struct GramBwdParams {
    int B, d, n;
    int n_tiles, j_tiles, N;          // N = MMA N (multiple of 16, <= 128)
    int nstages, stage_bytes;
    const float* gG;
    float* gV;
};

__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;     // next 32-element group along M/N
    d |= (uint64_t)(512u >> 4) << 32;                      // next atom of 4 k-rows (4 x 128 bytes)
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)1 << 61;                                // SWIZZLE_128B_BASE32B: the only MN-major tf32 layout
    return d;
}

__global__ void __launch_bounds__(kBwdThreads)
gram_tc_bwd_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmapX,
                   const __grid_constant__ CUtensorMap tmapY, const GramBwdParams P) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kMaxStages], split_bar[kMaxStages], empty_bar[kMaxStages], done_bar;
    __shared__ uint32_t tmem_base_smem;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // work item: j-tile fastest so CTAs sharing a V tile run together (L2 reuse)
    int w = blockIdx.x;
    const int jt = w % P.j_tiles; w /= P.j_tiles;
    const int nt = w % P.n_tiles;
    const int b = w / P.n_tiles;
    const int d = P.d, N = P.N;
    const int n0 = nt * kTile, j0 = jt * kTile;
    const int nk = (d + kKc - 1) / kKc;
    const int nst = P.nstages;

    if (threadIdx.x == 0) {
        for (int s = 0; s < nst; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&split_bar[s], 128);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(&done_bar, 1);
        mbar_fence_init();
        prefetch_tmap(&tmap);
        prefetch_tmap(&tmapX);
        prefetch_tmap(&tmapY);
    }
    if (warp == 1) tmem_alloc<kBwdTmemCols>(&tmem_base_smem);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    if (warp == 0) {
        if (lane == 0) {
            for (int kc = 0; kc < nk; ++kc) {
                const int s = kc % nst;
                const uint32_t ph = (uint32_t)(kc / nst) & 1u;
                mbar_wait(&empty_bar[s], ph ^ 1u);
                uint8_t* st = smem + (size_t)s * P.stage_bytes;
                mbar_arrive_expect_tx(&full_bar[s], (uint32_t)(kTileBytes + 256 * N));
#pragma unroll
                for (int g = 0; g < 4; ++g)       // four 32-wide n groups of the 128-wide tile
                    tma_load_2d(st + g * 4096, &tmap, &full_bar[s], n0 + g * 32, b * d + kc * kKc);
                // X = g_G[j0.., kc*32..] (K-major, 128B swizzle) and Y = g_G[kc*32.., j0..] (plain rows)
                tma_load_2d(st + 2 * kTileBytes, &tmapX, &full_bar[s], kc * kKc, b * d + j0);
                tma_load_2d(st + 2 * kTileBytes + 2 * N * 128, &tmapY, &full_bar[s], j0, b * d + kc * kKc);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_tf32(kTile, N) | (1u << 15);     // A is MN-major
            for (int kc = 0; kc < nk; ++kc) {
                const int s = kc % nst;
                const uint32_t ph = (uint32_t)(kc / nst) & 1u;
                mbar_wait(&split_bar[s], ph);
                tc_fence_after();
                const uint32_t a_hi = smem_u32(smem + (size_t)s * P.stage_bytes);
                const uint32_t a_lo = a_hi + kTileBytes;
                const uint32_t b_hi = a_lo + kTileBytes;
                const uint32_t b_lo = b_hi + (uint32_t)N * 128u;
                const int ksteps = min(kKc, d - kc * kKc + 7) / 8;
                for (int k = 0; k < ksteps; ++k) {
                    const uint64_t dah = umma_desc_sw128_mn(a_hi + k * 1024, 4096), dal = umma_desc_sw128_mn(a_lo + k * 1024, 4096);
                    const uint64_t dbh = umma_desc_sw128(b_hi + k * 32), dbl = umma_desc_sw128(b_lo + k * 32);
                    umma_tf32(tmem_base, dal, dbh, idesc, (kc | k) != 0);
                    umma_tf32(tmem_base, dah, dbl, idesc, 1);
                    umma_tf32(tmem_base, dah, dbh, idesc, 1);
                }
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&done_bar);
        }
    } else {
        const int t = threadIdx.x - 64;                      // 0..127
        for (int kc = 0; kc < nk; ++kc) {
            const int s = kc % nst;
            const uint32_t ph = (uint32_t)(kc / nst) & 1u;
            uint8_t* st = smem + (size_t)s * P.stage_bytes;
            mbar_wait(&full_bar[s], ph);
            // ---- B = X + Y^T (S = g_G + g_G^T), hi in place of X, lo beside it; k >= d masked
            {
                uint8_t* bh = st + 2 * kTileBytes;
                uint8_t* bl = bh + (size_t)N * 128;
                const float* Y = reinterpret_cast<const float*>(bl + (size_t)N * 128);
                const int kvalid = d - kc * kKc;             // rows of Y that belong to this sample
                for (int it = t; it < N * 8; it += 128) {
                    const int c = it / N, j = it - c * N;    // lanes run along j: conflict-free everywhere
                    const uint32_t off = (uint32_t)j * 128u + (((uint32_t)c ^ ((uint32_t)j & 7u)) << 4);
                    float4 x = *reinterpret_cast<const float4*>(bh + off);
                    const int k = 4 * c;
                    x.x = (k + 0 < kvalid) ? x.x + Y[(k + 0) * N + j] : 0.f;
                    x.y = (k + 1 < kvalid) ? x.y + Y[(k + 1) * N + j] : 0.f;
                    x.z = (k + 2 < kvalid) ? x.z + Y[(k + 2) * N + j] : 0.f;
                    x.w = (k + 3 < kvalid) ? x.w + Y[(k + 3) * N + j] : 0.f;
                    float4 h, l;
                    split_tf32(x.x, h.x, l.x); split_tf32(x.y, h.y, l.y);
                    split_tf32(x.z, h.z, l.z); split_tf32(x.w, h.w, l.w);
                    *reinterpret_cast<float4*>(bh + off) = h;
                    *reinterpret_cast<float4*>(bl + off) = l;
                }
            }
            // ---- A: split the landed V tile in place
            for (int i = t; i < kTileBytes / 16; i += 128) {
                float4* raw = reinterpret_cast<float4*>(st) + i;
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
        // ---- epilogue: lane = n, column = j
        mbar_wait(&done_bar, 0);
        tc_fence_after();
        const int q = warp & 3;
        const int nn = n0 + q * 32 + lane;
        float* Ob = P.gV + (size_t)b * d * P.n;
        for (int c0 = 0; c0 < N; c0 += 32) {
            float v[32];
            tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
            if (nn < P.n) {
#pragma unroll
                for (int u = 0; u < 32; ++u) {
                    const int j = j0 + c0 + u;
                    if (c0 + u < N && j < d) Ob[(size_t)j * P.n + nn] = v[u];
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kBwdTmemCols>(tmem_base);
    }
}

bool gram_tc_bwd_usable(int B, int d, int n, const float* V, const float* gG, const float* gV) {
    return B > 0 && d >= 16 && (d & 3) == 0 && n >= kKc && (n & 3) == 0 && aligned16(V) && aligned16(gG) && gV &&
           (long long)B * d < (1LL << 31);
}

int gram_tc_bwd_launch(int B, int d, int n, const float* V, const float* gG, float* gV, cudaStream_t stream) {
    GramBwdParams P{};
    P.B = B; P.d = d; P.n = n; P.gG = gG; P.gV = gV;
    P.n_tiles = (int)ceil_div(n, kTile);
    P.j_tiles = (int)ceil_div(d, kTile);
    P.N = d >= kTile ? kTile : (int)ceil_div(d, 16) * 16;
    P.stage_bytes = 2 * kTileBytes + 3 * P.N * 128;          // A hi/lo, B hi/lo, Y
    P.stage_bytes = (int)ceil_div(P.stage_bytes, 1024) * 1024;
    P.nstages = (int)min((long long)4, (190LL * 1024) / P.stage_bytes);
    const long long grid = (long long)B * P.n_tiles * P.j_tiles;
    CMF_REQUIRE(grid < (1LL << 31), "gram_tc_bwd: grid too large");
    CUtensorMap tmap;
    int rc = make_tmap_2d_f32(&tmap, V, (uint64_t)B * d, (uint64_t)n, (uint64_t)n, 32, 1);
    if (rc != 0) return rc;
    CUtensorMap tmapX, tmapY;
    rc = make_tmap_2d_f32(&tmapX, gG, (uint64_t)B * d, (uint64_t)d, (uint64_t)d, (uint32_t)P.N, 0);
    if (rc != 0) return rc;
    rc = make_tmap_2d_plain_f32(&tmapY, gG, (uint64_t)B * d, (uint64_t)d, (uint64_t)d, (uint32_t)P.N, 32);
    if (rc != 0) return rc;
    const size_t smem = (size_t)P.nstages * P.stage_bytes + 1024;
    static DeviceOnce attr_set;
    if (!attr_set.done()) {
        cudaError_t e = cudaFuncSetAttribute(gram_tc_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) { set_error("gram_tc_bwd attr: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set.mark();
    }
    gram_tc_bwd_kernel<<<(unsigned)grid, kBwdThreads, smem, stream>>>(tmap, tmapX, tmapY, P);
    CMF_CHECK_LAUNCH();
    return 0;
}

}  // namespace cmf
