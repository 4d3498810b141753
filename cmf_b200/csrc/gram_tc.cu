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
// Pipeline (512 threads): warp 0 = TMA producer (one lane; 128-byte-swizzled boxes of 32 k-columns),
// warp 1 = TMEM allocation + MMA issue (one lane), warps 2..7 = split the landed tile in shared
// memory (two groups on alternate stages; lo beside the raw tile, which is hi), warps 8..15 = drain (see "Accumulation precision") + epilogue.
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
constexpr int kLoDepth = 4;           // deepest lo ring (split -> MMA); deep enough that a split does not wait for the MMA two stages back
constexpr int kSplitGroup = 3;        // warps per split group; two groups take the even / odd stages
constexpr int kSplitWarps = 2 * kSplitGroup, kDrainWarps = 8;   // 16 warps: 128 registers each for the drain warps
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
    int nlo;        // lo ring depth (<= kLoDepth)
    int raw_bytes;  // bytes of one raw stage (one or two 16 KB tiles)
    int box_rows;
    int accumulate;
    int ablate;     // timing experiments (CMF_GRAM_FWD_ABLATE): 1 no MMA, 2 no split
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
    __shared__ __align__(8) uint64_t split_bar[kLoDepth], lo_free[kLoDepth], hh_full[2], hh_free[2], done_bar;
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
    const int nst = P.nstages, nlo = P.nlo;
    uint8_t* lo_ring = smem + (size_t)nst * P.raw_bytes;      // nlo x raw_bytes

    if (threadIdx.x == 0) {
        for (int s = 0; s < nst; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int i = 0; i < P.nlo; ++i) {
            mbar_init(&split_bar[i], 32 * kSplitGroup);
            mbar_init(&lo_free[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
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
                const int s = kc % nst, l = kc % nlo;
                const int g = kc / kDrainEvery, r = kc - g * kDrainEvery, pp = g & 1;
                mbar_wait(&split_bar[l], (uint32_t)(kc / nlo) & 1u);
                if (r == 0) mbar_wait(&hh_free[pp], ((uint32_t)(g >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t a_hi = smem_u32(smem + (size_t)s * P.raw_bytes);
                const uint32_t a_lo = smem_u32(lo_ring + (size_t)l * P.raw_bytes);
                const uint32_t b_hi = same ? a_hi : a_hi + kTileBytes;
                const uint32_t b_lo = same ? a_lo : a_lo + kTileBytes;
                const uint32_t t_hh = tmem_base + 128u + 128u * (uint32_t)pp;
#pragma unroll
                for (int k = 0; k < kKc / 8; ++k)            // UMMA K = 8 for tf32 (32 bytes)
                    if (!(P.ablate & 1)) umma_tf32(t_hh, umma_desc_sw128(a_hi + k * 32), umma_desc_sw128(b_hi + k * 32), idesc, (r | k) != 0);
                if (r == kDrainEvery - 1 || kc == nk - 1) umma_commit(&hh_full[pp]);   // drain warps may read hh[pp]
#pragma unroll
                for (int k = 0; k < ((P.ablate & 1) ? 0 : kKc / 8); ++k) {
                    umma_tf32(tmem_base, umma_desc_sw128(a_lo + k * 32), umma_desc_sw128(b_hi + k * 32), idesc,
                              (kc | k) != 0);
                    // both operands are the same rows: hi lo^T is the transpose of lo hi^T and is added in the epilogue
                    // (the kernel is bound by the MMAs' shared-memory operand reads, 8 KB each)
                    if (!same) umma_tf32(tmem_base, umma_desc_sw128(a_hi + k * 32), umma_desc_sw128(b_lo + k * 32), idesc, 1);
                }
                umma_commit(&empty_bar[s]);                  // raw slot and lo slot free once these retire
                umma_commit(&lo_free[l]);
            }
            umma_commit(&done_bar);                          // cross accumulator complete
        }
    } else if (warp < 2 + kSplitWarps) {
        // ===== splitter warps: lo = v - trunc(v) into the lo ring.  hi is the landed tile itself: kind::tf32 reads the top
        // 19 bits of each word (G is bit-identical with and without clearing the low 13 bits in shared memory).
        // A stage's split is a latency chain for one warp per scheduler (LDS -> ALU -> STS -> proxy fence -> arrive,
        // ~1200 cycles measured): two groups work on alternate stages, group g owns slot g of the lo ring
        const int t = (threadIdx.x - 64) % (32 * kSplitGroup);   // thread within the group
        const int nvec = (same ? 1 : 2) * (kTileBytes / 16); // float4 per stage of raw data
        for (int kc = (warp - 2) / kSplitGroup; kc < nk; kc += 2) {
            const int s = kc % nst, l = kc % nlo;
            mbar_wait(&full_bar[s], (uint32_t)(kc / nst) & 1u);
            mbar_wait(&lo_free[l], ((uint32_t)(kc / nlo) & 1u) ^ 1u);
            float4* raw = reinterpret_cast<float4*>(smem + (size_t)s * P.raw_bytes);
            float4* lo = reinterpret_cast<float4*>(lo_ring + (size_t)l * P.raw_bytes);
#pragma unroll 4
            for (int i = t; i < ((P.ablate & 2) ? 0 : nvec); i += 32 * kSplitGroup) {
                const float4 v = raw[i];
                float4 h, w;
                split_tf32(v.x, h.x, w.x); split_tf32(v.y, h.y, w.y);
                split_tf32(v.z, h.z, w.z); split_tf32(v.w, h.w, w.w);
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
            if (needed && !(P.ablate & 8)) {
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
                for (int u = 0; u < 32; ++u)     // same rows: T = H / 2 + C, G = T + T^T below
                    Gs[i * kGsLd + c0 + u] = same ? fmaf(acc[ch][u], 0.5f, v[u]) : acc[ch][u] + v[u];
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
                const float g = Gs[(s * d + li) * kGsLd + s * d + j] + Gs[(s * d + j) * kGsLd + s * d + li];
                Gb[e] = accm ? Gb[e] + g : g;
            }
        } else {
            float* Gb = G + (size_t)b0 * d * d;
            for (int e = t; e < kTile * kTile; e += 32 * kDrainWarps) {
                const int r = e >> 7, c = e & 127;
                const int gi = tr * kTile + r, gj = tc * kTile + c;
                if (gi < d && gj < d) {
                    const float g = tr != tc ? Gs[r * kGsLd + c] : Gs[r * kGsLd + c] + Gs[c * kGsLd + r];
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
// Panel form of the forward kernel for d <= 128: persistent, A operand in tensor memory.
// The kernel above is bound by shared-memory wavefronts (per K = 32 stage: 64 KB of SS-mode operand reads, the split's
// 16 KB read + 16 KB write, TMA's 16 KB) and pays ~10 us of prologue + epilogue per wave of 148 CTAs.  Here
//   * a split thread owns one row of the panel (= one TMEM lane): it reads its 128 bytes of the landed, swizzled tile,
//     and writes hi | lo into a staging slot of tensor memory (tcgen05.st); the MMAs take A from there and only B (the
//     landed tile itself = hi, 4 KB per MMA) from shared memory: hi hi^T into the ping-pong accumulator, lo hi^T into
//     the cross accumulator (the transpose is added in the epilogue) -- 48 KB of shared-memory traffic per stage;
//   * one CTA per SM walks the panels blockIdx.x, blockIdx.x + gridDim.x, ...: all rings keep rolling (phases are
//     functions of the global stage / group / panel counters), the drain warps write panel i (staging tile of their
//     own, coalesced rows) while the producer, the split warps and the tensor core are up to two drain groups into
//     panel i + 1.
// ------------------------------------------------------------------------------------------------
constexpr int kPanelSplitGroups = 1;                      // groups of 4 split warps (one TMEM lane quarter each) on alternate stages
constexpr int kPanelIssuer2 = 2 + 4 * kPanelSplitGroups + kDrainWarps;   // second MMA issuer: the last warp
constexpr int kPanelThreads = 32 * (kPanelIssuer2 + 1);   // TMA, MMA (hi hi^T), split warps, 8 drain warps, MMA (lo hi^T)
constexpr uint32_t kPanelStg = 384;                       // TMEM: [0,128) cross, [128,384) hi hi^T ping-pong, [384,512) 2 x (32 hi | 32 lo)

struct GramPanelParams {
    int B, d, n, n2, S, npanels, nst, box_rows, accumulate;
    float* G;
};

__global__ void __launch_bounds__(kPanelThreads, 1)
gram_tc_fwd_panel_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap2,
                         const GramPanelParams P) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kMaxStages], empty_bar[kMaxStages];
    __shared__ __align__(8) uint64_t a_full[2], a_free[2], hh_full[2], hh_free[2], cross_full, cross_free;
    __shared__ uint32_t tmem_base_smem;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int d = P.d, nst = P.nst;
    const int nk1 = (P.n + kKc - 1) / kKc;
    const int nk = nk1 + (P.n2 + kKc - 1) / kKc;
    const int ngroups = (nk + kDrainEvery - 1) / kDrainEvery;
    const int nit = (P.npanels - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;    // panels of this CTA (>= 1)
    float* Gs = reinterpret_cast<float*>(smem + (size_t)nst * kTileBytes);                  // [128][kGsLd] staging of the epilogue

    if (threadIdx.x == 0) {
        for (int s = 0; s < nst; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 2);                // both issuers' MMAs on the stage have retired
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&a_full[i], 128);
            mbar_init(&a_free[i], 2);
            mbar_init(&hh_full[i], 1);
            mbar_init(&hh_free[i], 32 * kDrainWarps);
        }
        mbar_init(&cross_full, 1);
        mbar_init(&cross_free, 32 * kDrainWarps);
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
            const uint32_t tx = (uint32_t)P.box_rows * kKc * 4;
            int gs = 0;
            for (int it = 0; it < nit; ++it) {
                const int row0 = ((int)blockIdx.x + it * (int)gridDim.x) * P.S * d;
                for (int kc = 0; kc < nk; ++kc, ++gs) {
                    const int s = gs % nst;
                    mbar_wait(&empty_bar[s], (((uint32_t)(gs / nst)) & 1u) ^ 1u);
                    mbar_arrive_expect_tx(&full_bar[s], tx);
                    const CUtensorMap* tm = kc < nk1 ? &tmap : &tmap2;
                    tma_load_2d(smem + (size_t)s * kTileBytes, tm, &full_bar[s], (kc < nk1 ? kc : kc - nk1) * kKc, row0);
                }
            }
        }
    } else if (warp == 1 || warp == kPanelIssuer2) {
        // ===== MMA issuers: warp 1 the hi hi^T products (ping-pong accumulator), the last warp the lo hi^T products (cross
        // accumulator).  One thread issues a tcgen05.mma per ~80 cycles; an M = N = 128 MMA occupies the tensor pipe for 64.
        if (lane == 0) {
            const bool hh = warp == 1;
            const uint32_t idesc = umma_idesc_tf32(kTile, kTile);
            int gs = 0, gg = 0;
            for (int it = 0; it < nit; ++it) {
                if (!hh && it > 0) mbar_wait(&cross_free, ((uint32_t)(it - 1)) & 1u);   // the drain warps have read the cross accumulator
                for (int kc = 0; kc < nk; ++kc, ++gs) {
                    const int s = gs % nst, ab = gs & 1;
                    const int g = kc / kDrainEvery, r = kc - g * kDrainEvery;
                    const int pp = (gg + g) & 1;
                    mbar_wait(&a_full[ab], ((uint32_t)(gs >> 1)) & 1u);              // implies the tile has landed
                    if (hh && r == 0) mbar_wait(&hh_free[pp], (((uint32_t)((gg + g) >> 1)) & 1u) ^ 1u);
                    tc_fence_after();
                    const uint64_t db = umma_desc_sw128(smem_u32(smem + (size_t)s * kTileBytes));
                    const uint32_t a_hi = tmem_base + kPanelStg + 64u * (uint32_t)ab, a_lo = a_hi + 32u;
                    if (hh) {
                        const uint32_t t_hh = tmem_base + 128u + 128u * (uint32_t)pp;
#pragma unroll
                        for (int k = 0; k < kKc / 8; ++k) umma_tf32_ts(t_hh, a_hi + k * 8, db + 2 * k, idesc, (r | k) != 0);
                        if (r == kDrainEvery - 1 || kc == nk - 1) umma_commit(&hh_full[pp]);
                    } else {
#pragma unroll
                        for (int k = 0; k < kKc / 8; ++k) umma_tf32_ts(tmem_base, a_lo + k * 8, db + 2 * k, idesc, (kc | k) != 0);
                    }
                    umma_commit(&empty_bar[s]);
                    umma_commit(&a_free[ab]);
                }
                if (!hh) umma_commit(&cross_full);
                gg += ngroups;
            }
        }
    } else if (warp < 2 + 4 * kPanelSplitGroups) {
        // ===== split warps: thread = panel row = TMEM lane; group g takes the stages gs = g mod kPanelSplitGroups
        const int grp = (warp - 2) >> 2, q = warp & 3;
        const int row = q * 32 + lane;
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        const int total = nit * nk;
        for (int gs = grp; gs < total; gs += kPanelSplitGroups) {
            const int s = gs % nst, ab = gs & 1;
            mbar_wait(&full_bar[s], ((uint32_t)(gs / nst)) & 1u);
            const uint8_t* rp = smem + (size_t)s * kTileBytes + row * 128;
            // two halves of 16 k (register budget: 96 at 576 threads); chunk c of the row sits at position c ^ (row & 7)
            float hi[16], lo[16];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float4 v = *reinterpret_cast<const float4*>(rp + ((c ^ (row & 7)) << 4));
                split_tf32(v.x, hi[4 * c], lo[4 * c]);         split_tf32(v.y, hi[4 * c + 1], lo[4 * c + 1]);
                split_tf32(v.z, hi[4 * c + 2], lo[4 * c + 2]); split_tf32(v.w, hi[4 * c + 3], lo[4 * c + 3]);
            }
            mbar_wait(&a_free[ab], (((uint32_t)(gs >> 1)) & 1u) ^ 1u);
            tc_fence_after();
            const uint32_t stg = tmem_base + kPanelStg + 64u * (uint32_t)ab + lane_addr;
            tmem_st_32x16(stg, hi);
            tmem_st_32x16(stg + 32u, lo);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const float4 v = *reinterpret_cast<const float4*>(rp + (((c + 4) ^ (row & 7)) << 4));
                split_tf32(v.x, hi[4 * c], lo[4 * c]);         split_tf32(v.y, hi[4 * c + 1], lo[4 * c + 1]);
                split_tf32(v.z, hi[4 * c + 2], lo[4 * c + 2]); split_tf32(v.w, hi[4 * c + 3], lo[4 * c + 3]);
            }
            tmem_st_32x16(stg + 16u, hi);
            tmem_st_32x16(stg + 48u, lo);
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(&a_full[ab]);
        }
    } else if (warp < kPanelIssuer2) {
        // ===== drain warps: FP32 round-to-nearest accumulation of hi hi^T, then the epilogue of the panel
        const int dw = warp - (2 + 4 * kPanelSplitGroups);   // 0..7
        const int q = warp & 3;
        const int half = dw >> 2;                            // column half [64 half, 64 half + 64)
        const int i = q * 32 + lane;
        bool needed;
        {
            const int r0 = q * 32, r1 = q * 32 + 31;
            const int c_lo = (r0 / d) * d, c_hi = min((r1 / d) * d + d, P.S * d);
            needed = (r0 < P.S * d) && (half * 64 < c_hi) && (half * 64 + 64 > c_lo);
        }
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        const int t = threadIdx.x - 32 * (2 + 4 * kPanelSplitGroups);   // 0..255
        int gg = 0;
        for (int it = 0; it < nit; ++it) {
            float acc[2][32];
#pragma unroll
            for (int ch = 0; ch < 2; ++ch)
#pragma unroll
                for (int u = 0; u < 32; ++u) acc[ch][u] = 0.f;
            for (int g = 0; g < ngroups; ++g) {
                const int pp = (gg + g) & 1;
                mbar_wait(&hh_full[pp], ((uint32_t)((gg + g) >> 1)) & 1u);
                tc_fence_after();
                if (needed) {
                    const uint32_t t_hh = tmem_base + 128u + 128u * (uint32_t)pp + (uint32_t)(half * 64);
#pragma unroll
                    for (int ch = 0; ch < 4; ++ch) {
                        float v[16];
                        tmem_ld_32x16(t_hh + lane_addr + (uint32_t)(ch * 16), v);
#pragma unroll
                        for (int u = 0; u < 16; ++u) acc[ch >> 1][(ch & 1) * 16 + u] += v[u];
                    }
                }
                tc_fence_before();
                mbar_arrive(&hh_free[pp]);
            }
            gg += ngroups;
            mbar_wait(&cross_full, (uint32_t)it & 1u);
            tc_fence_after();
            asm volatile("bar.sync 1, %0;" ::"n"(32 * kDrainWarps) : "memory");     // the previous panel's rows have left Gs
            if (needed) {
#pragma unroll
                for (int ch = 0; ch < 4; ++ch) {
                    const int c0 = half * 64 + ch * 16;
                    float v[16];
                    tmem_ld_32x16(tmem_base + lane_addr + (uint32_t)c0, v);
#pragma unroll
                    for (int u = 0; u < 16; ++u) Gs[i * kGsLd + c0 + u] = fmaf(acc[ch >> 1][(ch & 1) * 16 + u], 0.5f, v[u]);   // T = H / 2 + C
                }
            }
            tc_fence_before();
            mbar_arrive(&cross_free);
            asm volatile("bar.sync 1, %0;" ::"n"(32 * kDrainWarps) : "memory");
            const int b0 = ((int)blockIdx.x + it * (int)gridDim.x) * P.S;
            const int dd = d * d;
            const int total = min(P.S, P.B - b0) * dd;
            float* Gb = P.G + (size_t)b0 * dd;               // the panel's samples are contiguous in G
            for (int e = t; e < total; e += 32 * kDrainWarps) {
                const int s = e / dd, rem = e - s * dd;
                const int li = rem / d, j = rem - li * d;
                const float gv = Gs[(s * d + li) * kGsLd + s * d + j] + Gs[(s * d + j) * kGsLd + s * d + li];   // G = T + T^T
                Gb[e] = P.accumulate ? Gb[e] + gv : gv;
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
    static const int ablate_env = [] { const char* e = getenv("CMF_GRAM_FWD_ABLATE"); return e ? atoi(e) : 0; }();
    P.ablate = ablate_env;
    static const bool panel_off = [] { const char* e = getenv("CMF_GRAM_FWD_PANEL"); return e && e[0] == '0'; }();
    if (d <= kTile && !panel_off) {
        GramPanelParams Q{};
        Q.B = B; Q.d = d; Q.n = n; Q.n2 = n2; Q.G = G; Q.accumulate = accumulate;
        Q.S = kTile / d; Q.npanels = (int)ceil_div(B, Q.S); Q.box_rows = Q.S * d; Q.nst = 8;
        CUtensorMap tm1, tm2;
        int rc1 = make_tmap_2d_f32(&tm1, V, (uint64_t)B * d, (uint64_t)n, (uint64_t)n, (uint32_t)Q.box_rows, 0);
        if (rc1 != 0) return rc1;
        tm2 = tm1;
        if (n2 > 0) {
            rc1 = make_tmap_2d_f32(&tm2, V2, (uint64_t)B * d, (uint64_t)n2, (uint64_t)n2, (uint32_t)Q.box_rows, 0);
            if (rc1 != 0) return rc1;
        }
        const size_t smem_p = (size_t)Q.nst * kTileBytes + (size_t)kTile * kGsLd * 4 + 1024;
        static DeviceOnce attr_set_p;
        if (!attr_set_p.done()) {
            cudaError_t e = cudaFuncSetAttribute(gram_tc_fwd_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
            if (e != cudaSuccess) { set_error("gram_tc attr: %s", cudaGetErrorString(e)); return (int)e; }
            attr_set_p.mark();
        }
        const unsigned ctas = (unsigned)min((long long)Q.npanels, (long long)kNumSMs);
        gram_tc_fwd_panel_kernel<<<ctas, kPanelThreads, smem_p, stream>>>(tm1, tm2, Q);
        CMF_CHECK_LAUNCH();
        return 0;
    }
    long long grid;
    if (d <= kTile) {
        P.S = kTile / d; P.T = 1; P.ntiles = 1;
        P.box_rows = P.S * d;
        P.nstages = 8; P.nlo = 2; P.raw_bytes = kTileBytes;
        grid = ceil_div(B, P.S);
    } else {
        P.S = 0; P.T = (int)ceil_div(d, kTile); P.ntiles = P.T * (P.T + 1) / 2;
        P.box_rows = kTile;
        P.nstages = 4; P.nlo = 2; P.raw_bytes = 2 * kTileBytes;
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
    const size_t smem = (size_t)(P.nstages + P.nlo) * P.raw_bytes + 1024;     // raw ring + lo ring
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

// ------------------------------------------------------------------------------------------------
// Persistent adjoint for d <= 128 (one j tile).  A work item is (sample, 128-wide n tile) with only ceil(d / 32) K
// stages, so a CTA per item (the kernel above, still used for d > 128) spends its life in the prologue (barrier init,
// TMEM allocation, first TMA round trip) and the epilogue: ~8 us per wave of 148 CTAs, 14 waves at B = 512, n = 392
// (110 us; the FP32 FMA kernel needs 59).  Here one CTA per SM walks a contiguous range of items:
//   * B operand: S = g_G + g_G^T (hi | lo, K-major 128B-swizzled chunks of 32 k) is built ONCE per sample into one of
//     nS buffers from X / Y tiles that land in a small ring of their own, and is reused by the sample's n tiles;
//   * A operand: the V tile goes to TENSOR MEMORY.  A first persistent form kept it in shared memory (MN-major SS-mode
//     MMAs, four swizzled 4 KB TMA boxes per stage): 43 us, paced by its MMAs -- 24 per item at ~86 cycles, each
//     reading 4 KB of A + 2 KB of B from shared memory, every V tile three times (lo S_hi, hi S_lo, hi S_hi).  Now, as
//     in the convolution kernels, a split thread owns one TMEM lane (= one n position): it reads its column of the
//     landed [32 k][128 n] tile (plain row-major: ONE 16 KB TMA box per stage), splits it and writes hi | lo into a
//     staging slot of tensor memory with tcgen05.st; the MMAs read only the S chunks (2 KB) from shared memory.  The
//     ring stage is free as soon as the split warps have read it; there is no lo ring and no MN-major descriptor;
//   * two MMA issuers (even / odd items = accumulator 0 / 1, each with its own pair of staging slots): one thread
//     issues a tcgen05.mma per ~50-80 cycles, more than the 32 cycles an M = 128, N = 64 MMA occupies the tensor pipe;
//   * the accumulator pair is double-buffered in TMEM: eight epilogue warps take item i while item i + 1 is split and
//     multiplied; for d <= 64 they stage the [d][128] tile in shared memory and ONE TMA store writes it (columns past
//     n are dropped by the tensor map), for larger d (no room for the staging pair) they store 128-byte lines.
// Measured (scripts/gram_probe.py, B = 512, d = 64, n = 392, us): CTA per item 110, FMA 59, SS-mode persistent 43,
// V in TMEM 41, + second issuer 36, + TMA store 30.  Not kept: cp.async.bulk.prefetch of the V boxes ahead of the ring.
// ------------------------------------------------------------------------------------------------
constexpr int kBwdSplitGroups = 2;
constexpr int kBwdEpiWarps = 8;        // two per TMEM lane quarter (even / odd 32-column chunks of the accumulator): the
                                       // epilogue is a latency chain per warp (4 warps 48 us, 8 warps 43 us, 16 x 16 columns 44.5 us)
constexpr int kBwdPThreads = 32 * (2 + 4 * kBwdSplitGroups + kBwdEpiWarps);   // TMA, MMA, 2 x 4 split warps, epilogue warps

struct GramBwdCParams {
    int B, d, n, n_tiles, N, nk;
    int nst;            // raw V stages (16 KB each; even, see the split warps)
    int nS, ngl;        // S buffers, landing slots for the X / Y tiles
    int s_bytes;        // one S buffer: 2 * nk * N * 128
    int ablate;         // timing experiments (CMF_GRAM_BWD_ABLATE): 1 no MMA, 2 no S build, 4 no V split, 8 no stores
    int tma_out;        // g_V leaves through a staged tile and TMA stores (d <= 64: there is room for the staging pair)
    float* gV;
};

constexpr int kBwdTsTmemCols = 512;   // [0, 256) two accumulators of <= 128 columns, [256, 512) per issuer 2 x (32 hi | 32 lo) staging
constexpr uint32_t kBwdTsStg = 256;
constexpr int kBwdTsIssuer2 = 2 + 4 * kBwdSplitGroups + kBwdEpiWarps;   // the second MMA issuer is the last warp
constexpr int kBwdTsThreads = kBwdPThreads + 32;

__global__ void __launch_bounds__(kBwdTsThreads, 1)
gram_tc_bwd_ts_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmapX,
                      const __grid_constant__ CUtensorMap tmapY, const __grid_constant__ CUtensorMap tmapO,
                      const GramBwdCParams P) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kMaxStages], empty_bar[kMaxStages];
    __shared__ __align__(8) uint64_t a_full[4], a_free[4], g_full[2], g_free[2], s_full[2], s_free[2];
    __shared__ __align__(8) uint64_t acc_full[2], acc_free[2];
    __shared__ uint32_t tmem_base_smem;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int d = P.d, N = P.N, nst = P.nst, nk = P.nk, nS = P.nS, ngl = P.ngl;
    const long long total = (long long)P.B * P.n_tiles;
    const int w_begin = (int)(total * blockIdx.x / gridDim.x), w_end = (int)(total * (blockIdx.x + 1) / gridDim.x);
    // shared-memory map: S buffers | X/Y landing slots | raw ring
    const int chunk_bytes = N * 128;
    uint8_t* S_base = smem;
    uint8_t* gl_base = S_base + (size_t)nS * P.s_bytes;
    uint8_t* out_base = gl_base + (size_t)ngl * 2 * chunk_bytes;          // tma_out: 2 x [d][128] output tiles
    uint8_t* raw_base = out_base + (P.tma_out ? (size_t)2 * d * kTile * 4 : 0);

    if (threadIdx.x == 0) {
        for (int s = 0; s < nst; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 128);
        }
        for (int i = 0; i < 4; ++i) {           // staging ring of issuer i / 2, slot i % 2
            mbar_init(&a_full[i], 128);
            mbar_init(&a_free[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&g_full[i], 1);
            mbar_init(&g_free[i], 128 * kBwdSplitGroups);
            mbar_init(&s_full[i], 128 * kBwdSplitGroups);
            mbar_init(&s_free[i], 2);
            mbar_init(&acc_full[i], 1);
            mbar_init(&acc_free[i], 32 * kBwdEpiWarps);
        }
        mbar_fence_init();
        prefetch_tmap(&tmap);
        prefetch_tmap(&tmapX);
        prefetch_tmap(&tmapY);
        if (P.tma_out) prefetch_tmap(&tmapO);
    }
    if (warp == 1) tmem_alloc<kBwdTsTmemCols>(&tmem_base_smem);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;

    if (warp == 0) {
        if (lane == 0) {
            int gs = 0, gg = 0, b_prev = -1;
            for (int w = w_begin; w < w_end; ++w) {
                const int b = w / P.n_tiles, n0 = (w - b * P.n_tiles) * kTile;
                if (b != b_prev) {                                   // the sample's g_G tiles, one landing slot per chunk
                    b_prev = b;
                    for (int kc = 0; kc < nk; ++kc, ++gg) {
                        const int gl = gg % ngl;
                        mbar_wait(&g_free[gl], (((uint32_t)(gg / ngl)) & 1u) ^ 1u);
                        uint8_t* dst = gl_base + (size_t)gl * 2 * chunk_bytes;
                        mbar_arrive_expect_tx(&g_full[gl], (uint32_t)(2 * chunk_bytes));
                        tma_load_2d(dst, &tmapX, &g_full[gl], kc * kKc, b * d);
                        tma_load_2d(dst + chunk_bytes, &tmapY, &g_full[gl], 0, b * d + kc * kKc);
                    }
                }
                for (int kc = 0; kc < nk; ++kc, ++gs) {
                    const int s = gs % nst;
                    mbar_wait(&empty_bar[s], (((uint32_t)(gs / nst)) & 1u) ^ 1u);
                    mbar_arrive_expect_tx(&full_bar[s], (uint32_t)kTileBytes);
                    tma_load_2d(raw_base + (size_t)s * kTileBytes, &tmap, &full_bar[s], n0, b * d + kc * kKc);
                }
            }
        }
    } else if (warp == 1 || warp == kBwdTsIssuer2) {
        // two issuers (even / odd items = accumulator 0 / 1): a single thread issues one tcgen05.mma per ~50-80 cycles,
        // more than the 32 cycles an M = 128, N = 64 MMA occupies the tensor pipe
        if (lane == 0) {
            const int me = warp == 1 ? 0 : 1;
            const uint32_t idesc = umma_idesc_tf32(kTile, N);
            int it = 0, sidx = -1, b_prev = -1;
            bool mine_in_sample = false;
            for (int w = w_begin; w < w_end; ++w, ++it) {
                const int b = w / P.n_tiles;
                if (b != b_prev) {
                    b_prev = b;
                    ++sidx;
                    mine_in_sample = false;
                }
                const int sb = sidx % nS;
                if ((it & 1) == me) {
                    if (!mine_in_sample) mbar_wait(&s_full[sb], ((uint32_t)(sidx / nS)) & 1u);
                    mine_in_sample = true;
                    const uint32_t s_addr = smem_u32(S_base + (size_t)sb * P.s_bytes);
                    mbar_wait(&acc_free[me], (((uint32_t)(it >> 1)) & 1u) ^ 1u);
                    const uint32_t t_acc = tmem_base + 128u * (uint32_t)me;
                    for (int kc = 0; kc < nk; ++kc) {
                        const int par = kc & 1, ab = 2 * me + par;       // staging slot; u = how often it was used before
                        const int u = (it >> 1) * ((nk + 1 - par) >> 1) + (kc >> 1);
                        mbar_wait(&a_full[ab], (uint32_t)u & 1u);
                        tc_fence_after();
                        const uint32_t a_hi = tmem_base + kBwdTsStg + 64u * (uint32_t)ab, a_lo = a_hi + 32u;
                        const uint64_t dbh = umma_desc_sw128(s_addr + (uint32_t)(kc * chunk_bytes));
                        const uint64_t dbl = umma_desc_sw128(s_addr + (uint32_t)((nk + kc) * chunk_bytes));
                        const int ksteps = min(kKc, d - kc * kKc + 7) / 8;
                        for (int k = 0; k < ((P.ablate & 1) ? 0 : ksteps); ++k) {   // +32 bytes along K = +2 in the address field
                            umma_tf32_ts(t_acc, a_hi + k * 8, dbl + 2 * k, idesc, (kc | k) != 0);
                            umma_tf32_ts(t_acc, a_lo + k * 8, dbh + 2 * k, idesc, 1);
                            umma_tf32_ts(t_acc, a_hi + k * 8, dbh + 2 * k, idesc, 1);
                        }
                        umma_commit(&a_free[ab]);
                    }
                    umma_commit(&acc_full[me]);
                }
                // end of the sample's tiles in this CTA: both issuers release the S buffer, each after its own MMAs
                if (w + 1 == w_end || (w + 1) / P.n_tiles != b) {
                    if (mine_in_sample) umma_commit(&s_free[sb]); else mbar_arrive(&s_free[sb]);
                }
            }
        }
    } else if (warp < 2 + 4 * kBwdSplitGroups) {
        const int ts = threadIdx.x - 64;                     // thread among all splitters (the S build is shared)
        const int grp = ts >> 7;                             // group g owns the stages gs = g mod 2
        const int q = warp & 3;                              // TMEM lane quarter of this warp
        const int nl = q * 32 + lane;                        // n position within the tile = TMEM lane
        const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
        int gs = 0, gg = 0, sidx = -1, b_prev = -1, it = 0;
        for (int w = w_begin; w < w_end; ++w, ++it) {
            const int b = w / P.n_tiles, n0 = (w - b * P.n_tiles) * kTile;
            const bool live = n0 + q * 32 < P.n;             // warp-uniform: this quarter holds data
            if (b != b_prev) {
                // S = g_G + g_G^T of the sample: hi | lo chunks in the operand layout; k >= d masked
                b_prev = b;
                ++sidx;
                const int sb = sidx % nS;
                mbar_wait(&s_free[sb], (((uint32_t)(sidx / nS)) & 1u) ^ 1u);
                uint8_t* Sb = S_base + (size_t)sb * P.s_bytes;
                for (int kc = 0; kc < nk; ++kc, ++gg) {
                    const int gl = gg % ngl;
                    mbar_wait(&g_full[gl], ((uint32_t)(gg / ngl)) & 1u);
                    const uint8_t* X = gl_base + (size_t)gl * 2 * chunk_bytes;
                    const float* Y = reinterpret_cast<const float*>(X + chunk_bytes);
                    uint8_t* bh = Sb + (size_t)kc * chunk_bytes;
                    uint8_t* bl = Sb + (size_t)(nk + kc) * chunk_bytes;
                    const int kvalid = d - kc * kKc;
                    for (int i = ts; i < ((P.ablate & 2) ? 0 : N * 8); i += 128 * kBwdSplitGroups) {
                        const int c = i / N, j = i - c * N;  // lanes run along j: conflict-free everywhere
                        const uint32_t off = (uint32_t)j * 128u + (((uint32_t)c ^ ((uint32_t)j & 7u)) << 4);
                        float4 x = *reinterpret_cast<const float4*>(X + off);
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
                    mbar_arrive(&g_free[gl]);
                }
                fence_proxy_async_smem();
                mbar_arrive(&s_full[sb]);
            }
            for (int kc = 0; kc < nk; ++kc, ++gs) {
                if ((gs & (kBwdSplitGroups - 1)) != grp) continue;
                const int s = gs % nst;
                // staging slot (issuer it & 1, parity of kc) and its use count: with an even ring depth a ring slot, and with
                // this slot rule a staging slot, is always served by the same split group, so every barrier is waited on in
                // phase order by one group (a waiter two phases ahead would pass on the parity of the phase before)
                const int par = kc & 1, ab = 2 * (it & 1) + par;
                const int u = (it >> 1) * ((nk + 1 - par) >> 1) + (kc >> 1);
                mbar_wait(&full_bar[s], ((uint32_t)(gs / nst)) & 1u);
                const float* raw = reinterpret_cast<const float*>(raw_base + (size_t)s * kTileBytes) + nl;   // [32 k][128 n]
                float hi[32], lo[32];
                if (live && !(P.ablate & 4)) {
#pragma unroll
                    for (int k = 0; k < 32; ++k) {
                        const float v = raw[k * kTile];
                        split_tf32(v, hi[k], lo[k]);
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < 32; ++k) hi[k] = lo[k] = 0.f;
                }
                mbar_arrive(&empty_bar[s]);                  // the values are in registers: the ring stage may be refilled
                mbar_wait(&a_free[ab], ((uint32_t)u & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t stg = tmem_base + kBwdTsStg + 64u * (uint32_t)ab + lane_addr;
                tmem_st_32x32(stg, hi);
                tmem_st_32x32(stg + 32u, lo);
                tmem_st_wait();
                tc_fence_before();
                mbar_arrive(&a_full[ab]);
            }
        }
    } else if (warp < kBwdTsIssuer2) {
        // epilogue warps (any four consecutive warps cover the four TMEM lane quarters): TMEM lane = n
        const int q = warp & 3;
        const int ew = (warp - (2 + 4 * kBwdSplitGroups)) >> 2;   // which of the quarter's warps
        const bool elected = warp == 2 + 4 * kBwdSplitGroups && lane == 0;
        int it = 0;
        for (int w = w_begin; w < w_end; ++w, ++it) {
            const int b = w / P.n_tiles, n0 = (w - b * P.n_tiles) * kTile;
            const int buf = it & 1;
            mbar_wait(&acc_full[buf], ((uint32_t)(it >> 1)) & 1u);
            tc_fence_after();
            const int nn = n0 + q * 32 + lane;
            const bool live = n0 + q * 32 < P.n;             // warp-uniform
            if (P.tma_out) {
                // accumulator -> registers -> staged [d][128] tile -> one TMA store (columns past n are dropped by the
                // tensor map): the warps are done with the item after 32 shared-memory stores instead of 32 global ones
                float v[32];
                const int c0 = 32 * ew;                       // d <= 64: one 32-column chunk per warp
                const bool has = live && c0 < N;
                if (has) tmem_ld_32x32(tmem_base + 128u * (uint32_t)buf + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
                tc_fence_before();
                mbar_arrive(&acc_free[buf]);                  // the accumulator is in registers
                float* stg = reinterpret_cast<float*>(out_base + (size_t)buf * d * kTile * 4) + q * 32 + lane;
                asm volatile("bar.sync 1, %0;" ::"n"(32 * kBwdEpiWarps) : "memory");   // the store two items back has read this tile
                if (has && !(P.ablate & 8)) {
#pragma unroll
                    for (int u = 0; u < 32; ++u)
                        if (c0 + u < d) stg[(c0 + u) * kTile] = v[u];
                }
                fence_proxy_async_smem();
                asm volatile("bar.sync 1, %0;" ::"n"(32 * kBwdEpiWarps) : "memory");
                if (elected) {
                    if (!(P.ablate & 8)) tma_store_2d(&tmapO, out_base + (size_t)buf * d * kTile * 4, n0, b * d);
                    bulk_commit();
                    bulk_wait_read<1>();                      // the other tile (next item's) is free again
                }
            } else {
                float* Ob = P.gV + (size_t)b * d * P.n;
                for (int c0 = 32 * ew; c0 < N; c0 += 32 * (kBwdEpiWarps / 4)) {
                    if (live) {
                        float v[32];
                        tmem_ld_32x32(tmem_base + 128u * (uint32_t)buf + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
                        if (nn < P.n && !(P.ablate & 8)) {
#pragma unroll
                            for (int u = 0; u < 32; ++u)
                                if (c0 + u < d) Ob[(size_t)(c0 + u) * P.n + nn] = v[u];
                        }
                    }
                }
                tc_fence_before();
                mbar_arrive(&acc_free[buf]);
            }
        }
        if (P.tma_out && elected) bulk_wait_all();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kBwdTsTmemCols>(tmem_base);
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
    static const bool persist_off = [] { const char* e = getenv("CMF_GRAM_BWD_PERSIST"); return e && e[0] == '0'; }();
    if (P.j_tiles == 1 && !persist_off) {
        static const int ablate_env = [] { const char* e = getenv("CMF_GRAM_BWD_ABLATE"); return e ? atoi(e) : 0; }();
        static const bool tma_out_off = [] { const char* e = getenv("CMF_GRAM_BWD_TMA_OUT"); return e && e[0] == '0'; }();
        GramBwdCParams C{};
        C.B = B; C.d = d; C.n = n; C.n_tiles = P.n_tiles; C.N = P.N; C.nk = (int)ceil_div(d, kKc);
        C.s_bytes = 2 * C.nk * C.N * 128;
        C.nS = d <= 64 ? 2 : 1;
        C.ngl = d <= 64 ? 2 : 1;
        C.ablate = ablate_env; C.gV = gV;
        C.tma_out = d <= 64 && !tma_out_off;
        CUtensorMap tmapV;                   // plain [32 k][128 n] boxes
        rc = make_tmap_2d_plain_f32(&tmapV, V, (uint64_t)B * d, (uint64_t)n, (uint64_t)n, kTile, kKc);
        if (rc != 0) return rc;
        CUtensorMap tmapO = tmapV;           // g_V as [d rows][128 n] boxes of one sample (never crosses into the next one)
        if (C.tma_out) {
            rc = make_tmap_2d_plain_f32(&tmapO, gV, (uint64_t)B * d, (uint64_t)n, (uint64_t)n, kTile, (uint32_t)d);
            if (rc != 0) return rc;
        }
        const long long fixed = (long long)C.nS * C.s_bytes + (long long)C.ngl * 2 * C.N * 128 + (C.tma_out ? 2LL * d * kTile * 4 : 0);
        C.nst = (int)min((long long)kMaxStages, (225LL * 1024 - fixed) / kTileBytes) & ~1;   // even: see the split warps
        CMF_REQUIRE(C.nst >= 2, "gram_tc_bwd: no room for the V ring (d=%d)", d);
        const size_t smem_ts = (size_t)fixed + (size_t)C.nst * kTileBytes + 1024;
        static DeviceOnce attr_set_ts;
        if (!attr_set_ts.done()) {
            cudaError_t e = cudaFuncSetAttribute(gram_tc_bwd_ts_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024);
            if (e != cudaSuccess) { set_error("gram_tc_bwd attr: %s", cudaGetErrorString(e)); return (int)e; }
            attr_set_ts.mark();
        }
        const unsigned ctas = (unsigned)min(grid, (long long)kNumSMs);
        gram_tc_bwd_ts_kernel<<<ctas, kBwdTsThreads, smem_ts, stream>>>(tmapV, tmapX, tmapY, tmapO, C);
        CMF_CHECK_LAUNCH();
        return 0;
    }
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
