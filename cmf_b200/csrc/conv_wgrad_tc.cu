// Weight gradient of the 3x3 / 64 -> 64 / pad 1 conditioner convolution on tcgen05 (3xTF32).
//
//     g_w[co, ci, tap] = sum_{n, p} g_y[n, co, p] * x[n, ci, p + dy*W + dx]        (p = flat pixel)
//
// The contraction runs over pixels, so pixels are the K dimension:
//   A (tensor memory, M = 128 rows) = TWO taps x 64 input channels of shifted x: row (a, ci) holds
//       x[ci, p + shift(2j+a)] for the 32 pixels of a sub-chunk.  Row workers (one thread per row)
//       read their 32 consecutive pixels from the halo tile TMA left in shared memory, zero the
//       pixels whose tap leaves the image row, split (round-to-nearest TF32 head + exact
//       remainder) and write both with tcgen05.st -- the same staging scheme as the forward kernel.
//   B (shared memory, N = 64 rows)  = g_y[co, p], K-major exactly as NCHW stores it: TMA drops
//       [64 co] x [32 px] boxes in the 128-byte swizzle; two splitter warps rewrite each box in
//       place into its TF32 head and write the remainder to a twin buffer (same swizzled
//       offsets, so no address arithmetic).
//   D (tensor memory) = 5 accumulators of 128 x 64: tap pair j -> g_w[:, :, 2j], g_w[:, :, 2j+1]
//       (the 10th "tap" is a zero row block).
// A CTA walks T consecutive 128-pixel tiles (default 14 = two 28x28 images) with all five
// accumulators resident in TMEM, then writes its partial g_w; a second kernel sums the partials
// in a fixed order with FP32 round-to-nearest adds.  T bounds the length of the
// truncating tensor-core accumulation: <= 672 MMAs per accumulator, a worst-case same-sign bias
// of ~3e-5 -- adequate for a gradient that goes straight to the optimiser (nothing downstream
// amplifies it, unlike the forward values that pass through G^-1).
#include "common.cuh"
#include "tc_common.cuh"

namespace cmf {
using namespace tc;

namespace {

constexpr int kC = 64;
constexpr int kTaps = 9;
constexpr int kPairs = 5;               // tap pairs (the last pair is tap 8 + a zero block)
constexpr int kTileM = 128;             // pixels per tile
constexpr int kSub = 4;                 // sub-chunks of 32 pixels (= K per staging)
constexpr int kStagings = kSub * kPairs;          // 20 per tile
constexpr int kStgBufs = 3;
constexpr int kBBytes = kC * 32 * 4;    // one [64 co][32 px] box: 8 KB
constexpr int kWgThreads = 32 * 13;     // TMA, MMA 0, 8 row workers, 2 splitters, MMA 1
constexpr int kMmaWarp1 = 12;
#ifndef CMF_WG_ISSUERS
#define CMF_WG_ISSUERS 2
#endif
constexpr int kIssuers = CMF_WG_ISSUERS;
constexpr int kAccCols = kPairs * kC;   // 320
constexpr int kWgTmemCols = 512;

__device__ __forceinline__ float lds_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
__device__ __forceinline__ float4 lds_v4(uint32_t addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ void sts_v4(uint32_t addr, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float tf32_head_rn(float v) {
    return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xFFFFE000u);
}
// Truncated head (one operation instead of two) for the x operand of the weight gradient: the
// remainder then carries 13 bits of which the tensor core keeps 11, a representation error of
// 2^-21 |x| per element instead of 2^-22 -- far below the truncating accumulate (1e-6 .. 1e-5 of
// the result), and nothing downstream amplifies a weight gradient.
__device__ __forceinline__ float tf32_head_trunc(float v) {
    return __uint_as_float(__float_as_uint(v) & 0xFFFFE000u);
}

struct WgradParams {
    int N, H, W, HW;
    int tiles_per_image;
    long long total_tiles;
    int tiles_per_cta;                  // tiles per chunk
    long long num_chunks;               // ceil(total_tiles / tiles_per_cta); CTA b walks chunks b, b + grid, ...
    int hl;
    float* partial;                     // [grid][9][64 ci][64 co]
};

// The last tile of an image is ragged (28x28: 16 of 128 pixels, 14x14: 68): its 32-pixel sub-chunks
// that lie entirely past the image are all-zero K-steps and are skipped -- 25 instead of 28
// sub-chunks per 28x28 image, 7 instead of 8 per 14x14 image.  Every role derives the same count.
__device__ __forceinline__ int live_subchunks(int hw, int p0) {
    const int n = (hw - p0 + 31) >> 5;
    return n < kSub ? n : kSub;
}

template <int kPitch, bool kDet>
__global__ void __launch_bounds__(kWgThreads, 1)
conv3x3_wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmx, const __grid_constant__ CUtensorMap tmg, WgradParams P) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ uint64_t x_full[2], g_full[2], tile_done[2];
    __shared__ uint64_t b_ready[kSub], qdone[kSub];
    __shared__ uint64_t st_full[kStgBufs], st_done[kStgBufs];
    __shared__ uint64_t acc_init;
    __shared__ uint32_t tmem_base_smem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    constexpr int kXBytes = kC * kPitch * 4;
    uint8_t* gyhi = smem;                                   // [2 tiles][4 sub-chunks][8 KB], swizzled boxes
    uint8_t* gylo = smem + 2 * kSub * kBBytes;              // [4][8 KB]
    uint8_t* xh = gylo + kSub * kBBytes;                    // [2][64 ci][kPitch px]

    // PERSISTENT: a CTA walks chunks of tiles_per_cta tiles.  The accumulators are re-zeroed per chunk
    // (the chunk bounds the truncating tensor-core accumulation) and the chunk's result is ADDED, in
    // FP32 round-to-nearest, to this CTA's own partial in global memory -- 148 x 147 KB stays in L2,
    // where one partial per chunk (1600 for the 28x28 tangent batch) cost 46 GB of DRAM traffic and
    // 9 ms per step in the reduction kernel.  Every mbarrier phase is a function of the running tile
    // counter gt (or staging counter ns), which all roles advance identically.
    auto chunk_tiles = [&](long long ch, long long& tile_begin) {
        tile_begin = ch * P.tiles_per_cta;
        return (int)min((long long)P.tiles_per_cta, P.total_tiles - tile_begin);
    };

    if (threadIdx.x == 0) {
        for (int b = 0; b < 2; ++b) { mbar_init(&x_full[b], 1); mbar_init(&g_full[b], 1); mbar_init(&tile_done[b], kIssuers); }
        for (int q = 0; q < kSub; ++q) { mbar_init(&b_ready[q], 2); mbar_init(&qdone[q], kIssuers); }
        for (int b = 0; b < kStgBufs; ++b) { mbar_init(&st_full[b], 4); mbar_init(&st_done[b], kIssuers); }
        mbar_init(&acc_init, 8);
        mbar_fence_init();
        prefetch_tmap(&tmx);
        prefetch_tmap(&tmg);
    }
    if (warp == 1) tmem_alloc<kWgTmemCols>(&tmem_base_smem);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;
    constexpr uint32_t kStgCol = kAccCols;                  // staging (buffer b) at columns 320 + 64 b: hi | lo

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int gt = 0;
            // x and g_y stream through once (1.3 GB per tangent-batch launch); the per-CTA partials
            // that are re-read between chunks should be what stays in L2
            const uint64_t pol = l2_policy_evict_first();
            for (long long ch = blockIdx.x; ch < P.num_chunks; ch += gridDim.x) {
                long long tile_begin;
                const int nt = chunk_tiles(ch, tile_begin);
                for (int i = 0; i < nt; ++i, ++gt) {
                    const int buf = gt & 1;
                    const long long T = tile_begin + i;
                    const int img = (int)(T / P.tiles_per_image), p0 = (int)(T % P.tiles_per_image) * kTileM;
                    if (gt >= 2) mbar_wait(&tile_done[buf], ((gt >> 1) - 1) & 1);
                    mbar_arrive_expect_tx(&x_full[buf], (uint32_t)kXBytes);
                    tma_load_2d_hint(xh + (size_t)buf * kXBytes, &tmx, &x_full[buf], p0 - P.hl, img * kC, pol);
                    mbar_arrive_expect_tx(&g_full[buf], (uint32_t)(kSub * kBBytes));
                    for (int q = 0; q < kSub; ++q)
                        tma_load_2d_hint(gyhi + (size_t)(buf * kSub + q) * kBBytes, &tmg, &g_full[buf], p0 + 32 * q, img * kC, pol);
                }
            }
        }
    } else if (warp == 1 || (warp == kMmaWarp1 && kIssuers == 2)) {
        // ------------------------------------------------------------------ MMA issuers
        // A single thread's issue path costs ~50 cycles per tcgen05.mma (240 MMAs per tile would
        // take 1.5x the tensor time), so TWO warps issue.  Both walk EVERY staging -- each sees every
        // phase of st_full / b_ready, which the parity waits require, and both commit st_done.
        //   default: issuer w takes the K-steps {2w, 2w+1} of every staging.  Both add into the same
        //     accumulator (zeroed by the row workers up front, so no MMA depends on being first); the
        //     relative order of the two MMA streams, hence the truncation pattern of the FP32
        //     accumulate, depends on timing: results are reproducible to 1-3 ulp, not bitwise.
        //   kDet: issuer w issues whole stagings, those of the accumulators j with j % 2 == w (tap
        //     pairs 0, 2, 4 / 1, 3).  Every accumulator sees ONE in-order MMA stream -- bitwise
        //     reproducible -- and the chunk's first MMA into it overwrites it (no zero-fill pass).
        //     A staging is released after 12 serially issued MMAs instead of 6: ~9 % slower.
        const int w = warp == 1 ? 0 : 1;
        constexpr uint32_t idesc = umma_idesc_tf32(kTileM, kC);
        const uint32_t gyhi_addr = smem_u32(gyhi), gylo_addr = smem_u32(gylo);
        int ns = 0, gt = 0, r = 0;
        for (long long ch = blockIdx.x; ch < P.num_chunks; ch += gridDim.x, ++r) {
        long long tile_begin;
        const int nt = chunk_tiles(ch, tile_begin);
        mbar_wait(&acc_init, r & 1);            // the row workers have read (and, !kDet, re-zeroed) the accumulators
        tc_fence_after();
        for (int i = 0; i < nt; ++i, ++gt) {
            const int buf = gt & 1;
            const int nq = live_subchunks(P.HW, (int)((tile_begin + i) % P.tiles_per_image) * kTileM);
            const int cnt = nq * kPairs;
            for (int n = 0; n < cnt; ++n, ++ns) {
                const int q = n / kPairs, j = n - q * kPairs;
                const int sb = ns % kStgBufs;
                const bool mine = !kDet || kIssuers == 1 || (j & 1) == w;
                if (j == 0) mbar_wait(&b_ready[q], gt & 1);
                mbar_wait(&st_full[sb], (ns / kStgBufs) & 1);
                tc_fence_after();
                if (elect_one()) {
                    const uint32_t acc = tmem_base + (uint32_t)(j * kC);
                    const uint32_t a_hi = tmem_base + kStgCol + (uint32_t)(sb * 64), a_lo = a_hi + 32;
                    const uint64_t dbh = umma_desc_sw128(gyhi_addr + (uint32_t)((buf * kSub + q) * kBBytes));
                    const uint64_t dbl = umma_desc_sw128(gylo_addr + (uint32_t)(q * kBBytes));
                    if (mine) {
                        constexpr int kPer = kDet ? 4 : 4 / kIssuers;
#pragma unroll
                        for (int kk = 0; kk < kPer; ++kk) {
                            const int k = kDet ? kk : w * kPer + kk;
                            // kDet: the chunk's first MMA into accumulator j overwrites it (q = 0 of the first tile is always live)
                            umma_tf32_ts(acc, a_hi + k * 8, dbl + 2 * k, idesc, kDet ? (i | q | k) != 0 : 1);
                            umma_tf32_ts(acc, a_lo + k * 8, dbh + 2 * k, idesc, 1);
                            umma_tf32_ts(acc, a_hi + k * 8, dbh + 2 * k, idesc, 1);
                        }
                    }
                    umma_commit(&st_done[sb]);
                    if (j == kPairs - 1) umma_commit(&qdone[q]);
                }
                __syncwarp();
            }
            // skipped sub-chunks: keep the b_ready / qdone phases in step (one per tile), and release
            // the tile's buffers only after the splitters are done rewriting ALL of its g_y boxes
            for (int q = nq; q < kSub; ++q) {
                mbar_wait(&b_ready[q], gt & 1);
                if (elect_one()) umma_commit(&qdone[q]);
                __syncwarp();
            }
            if (elect_one()) umma_commit(&tile_done[buf]);
            __syncwarp();
        }
        }
    } else if (warp == kMmaWarp1) {
        // second issuer slot unused (kIssuers == 1)
    } else if (warp >= 10) {
        // ------------------------------------------------------------------ g_y splitters (64 threads)
        const int tid = threadIdx.x - 320;
        const uint32_t gyhi_addr = smem_u32(gyhi), gylo_addr = smem_u32(gylo);
        int gt = 0;
        for (long long ch = blockIdx.x; ch < P.num_chunks; ch += gridDim.x) {
        long long tile_begin;
        const int nt = chunk_tiles(ch, tile_begin);
        for (int i = 0; i < nt; ++i, ++gt) {
            const int buf = gt & 1;
            mbar_wait(&g_full[buf], (gt >> 1) & 1);
            for (int q = 0; q < kSub; ++q) {
                if (gt >= 1) mbar_wait(&qdone[q], (gt - 1) & 1);      // remainder buffer q free again
                const uint32_t src = gyhi_addr + (uint32_t)((buf * kSub + q) * kBBytes), dst = gylo_addr + (uint32_t)(q * kBBytes);
#pragma unroll
                for (int it = 0; it < kBBytes / 16 / 64; ++it) {
                    const uint32_t off = (uint32_t)((tid + 64 * it) * 16);
                    const float4 v = lds_v4(src + off);
                    float4 h, l;
                    h.x = tf32_head_rn(v.x); h.y = tf32_head_rn(v.y); h.z = tf32_head_rn(v.z); h.w = tf32_head_rn(v.w);
                    l.x = v.x - h.x; l.y = v.y - h.y; l.z = v.z - h.z; l.w = v.w - h.w;
                    sts_v4(src + off, h);
                    sts_v4(dst + off, l);
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(&b_ready[q]);
            }
        }
        }
    } else {
        // ------------------------------------------------------------------ row workers (x operand)
        const int g = (warp - 2) >> 2;                    // this group stages n with n % 2 == g
        const int rb = warp & 3;
        const int m = rb * 32 + lane;
        const int a = m >> 6, ci = m & 63;
        const uint32_t lane_addr = (uint32_t)(rb * 32) << 16;
        const uint32_t xrow_addr = smem_u32(xh) + (uint32_t)(ci * kPitch * 4);
        int ns_base = 0, gt = 0, r = 0;                   // stagings / tiles / chunks before the current one
        for (long long ch = blockIdx.x; ch < P.num_chunks; ch += gridDim.x, ++r) {
        long long tile_begin;
        const int nt = chunk_tiles(ch, tile_begin);
        {   // the accumulators are free (their previous contents have been read by every worker warp);
            // shared-accumulator issue: zero the five accumulators (group g: columns [160 g, 160 g + 160)
            // of this warp's lane quarter)
            if (!kDet) {
                float zero[32];
#pragma unroll
                for (int k = 0; k < 32; ++k) zero[k] = 0.f;
#pragma unroll
                for (int c = 0; c < 5; ++c) tmem_st_32x32(tmem_base + lane_addr + (uint32_t)(g * 160 + c * 32), zero);
                tmem_st_wait();
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_init);
        }
        for (int i = 0; i < nt; ++i, ++gt) {
            const int buf = gt & 1;
            const long long T = tile_begin + i;
            const int p0 = (int)(T % P.tiles_per_image) * kTileM;
            const int cnt = live_subchunks(P.HW, p0) * kPairs;
            // bit k of ml[q] / mr[q]: pixel p0 + 32 q + k may use a tap with dx = -1 / +1
            uint32_t ml[kSub], mr[kSub];
#pragma unroll
            for (int q = 0; q < kSub; ++q) {
                const int xx = (p0 + 32 * q + lane) % P.W;
                ml[q] = __ballot_sync(0xffffffffu, xx > 0);
                mr[q] = __ballot_sync(0xffffffffu, xx + 1 < P.W);
            }
            mbar_wait(&x_full[buf], (gt >> 1) & 1);
            // group g stages the stagings whose GLOBAL index is congruent to g mod 2 (cnt may be odd)
            for (int n = (ns_base & 1) == g ? 0 : 1; n < cnt; n += 2) {
                const int ns = ns_base + n;
                const int q = n / kPairs, j = n - q * kPairs;
                const int tap = 2 * j + a;
                const int sb = ns % kStgBufs;
                float hi[32], lo[32];
                if (tap < kTaps) {
                    const int dy = tap / 3 - 1, dx = tap - (tap / 3) * 3 - 1;
                    // 32 consecutive pixels of row ci starting at `base`.  Rows are kPitch floats
                    // apart with kPitch/4 odd, so 128-bit loads are bank-conflict free across the 32
                    // channel lanes (32-bit loads would be 4-way conflicted); the window is loaded
                    // 16-byte aligned (9 vectors) and the misalignment base & 3 -- the same for every
                    // lane -- is resolved by a 4-way static dispatch.
                    const int base = P.hl + 32 * q + dy * P.W + dx;
                    const int off = base & 3;
                    const uint32_t src = xrow_addr + (uint32_t)((buf * kC * kPitch + (base - off)) * 4);
                    float wv[36];
#pragma unroll
                    for (int e = 0; e < 9; ++e) {
                        const float4 t4 = lds_v4(src + 16 * e);
                        wv[4 * e] = t4.x; wv[4 * e + 1] = t4.y; wv[4 * e + 2] = t4.z; wv[4 * e + 3] = t4.w;
                    }
                    uint32_t mlq = ml[0], mrq = mr[0];
#pragma unroll
                    for (int qq = 1; qq < kSub; ++qq) { if (q == qq) { mlq = ml[qq]; mrq = mr[qq]; } }
                    const uint32_t mb = dx < 0 ? mlq : (dx > 0 ? mrq : 0xffffffffu);
#define CMF_WG_SPLIT(OFF)                                                                         \
    _Pragma("unroll") for (int k = 0; k < 32; k += 2) {                                           \
        const float v0 = (mb & (1u << k)) ? wv[k + OFF] : 0.f;                                    \
        const float v1 = (mb & (2u << k)) ? wv[k + 1 + OFF] : 0.f;                                \
        hi[k] = tf32_head_trunc(v0);                                                                 \
        hi[k + 1] = tf32_head_trunc(v1);                                                             \
        fsub2(lo[k], lo[k + 1], v0, v1, hi[k], hi[k + 1]);                                        \
    }
                    // taps with dx = 0 (a third of the stagings; off is then P.W-dependent only) need no
                    // column mask: 3 of the 6 per-element operations go away
#define CMF_WG_SPLIT_NOMASK(OFF)                                                                  \
    _Pragma("unroll") for (int k = 0; k < 32; k += 2) {                                           \
        hi[k] = tf32_head_trunc(wv[k + OFF]);                                                        \
        hi[k + 1] = tf32_head_trunc(wv[k + 1 + OFF]);                                                \
        fsub2(lo[k], lo[k + 1], wv[k + OFF], wv[k + 1 + OFF], hi[k], hi[k + 1]);                  \
    }
                    if (dx == 0) {
                        if (off == 0) { CMF_WG_SPLIT_NOMASK(0) } else if (off == 1) { CMF_WG_SPLIT_NOMASK(1) }
                        else if (off == 2) { CMF_WG_SPLIT_NOMASK(2) } else { CMF_WG_SPLIT_NOMASK(3) }
                    } else {
                        if (off == 0) { CMF_WG_SPLIT(0) } else if (off == 1) { CMF_WG_SPLIT(1) }
                        else if (off == 2) { CMF_WG_SPLIT(2) } else { CMF_WG_SPLIT(3) }
                    }
#undef CMF_WG_SPLIT_NOMASK
#undef CMF_WG_SPLIT
                } else {
#pragma unroll
                    for (int k = 0; k < 32; ++k) { hi[k] = 0.f; lo[k] = 0.f; }
                }
                if (ns >= kStgBufs) mbar_wait(&st_done[sb], ((ns / kStgBufs) - 1) & 1);
                tc_fence_after();
                const uint32_t stg = tmem_base + kStgCol + (uint32_t)(sb * 64) + lane_addr;
                tmem_st_32x32(stg, hi);
                tmem_st_32x32(stg + 32, lo);
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&st_full[sb]);
            }
            ns_base += cnt;
        }
        // ------------------------------------------------------------------ epilogue: partial g_w
        {
            const int last = gt - 1;                      // every chunk has at least one tile
            mbar_wait(&tile_done[last & 1], (last >> 1) & 1);
            tc_fence_after();
            float* out = P.partial + (size_t)blockIdx.x * (kTaps * kC * kC);
            for (int j = (g == 0 ? 0 : 3); j < (g == 0 ? 3 : kPairs); ++j) {
                const int tap = 2 * j + a;
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    float v[32];
                    tmem_ld_32x32(tmem_base + lane_addr + (uint32_t)(j * kC + 32 * h), v);
                    if (tap < kTaps) {
                        float4* dst = reinterpret_cast<float4*>(out + ((size_t)tap * kC + ci) * kC + 32 * h);
                        if (r == 0) {
#pragma unroll
                            for (int e = 0; e < 8; ++e) dst[e] = make_float4(v[4 * e], v[4 * e + 1], v[4 * e + 2], v[4 * e + 3]);
                        } else {      // this thread wrote these words itself one chunk ago: plain read-add-write
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                const float4 o = dst[e];
                                dst[e] = make_float4(o.x + v[4 * e], o.y + v[4 * e + 1], o.z + v[4 * e + 2], o.w + v[4 * e + 3]);
                            }
                        }
                    }
                }
            }
            // the re-zeroing of the next chunk touches accumulator columns another worker warp of the
            // same lane quarter has just read: all 8 worker warps (256 threads) meet first
            tc_fence_before();
            asm volatile("bar.sync 1, 256;" ::: "memory");
            tc_fence_after();
        }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc<kWgTmemCols>(tmem_base);
    }
}

// g_w[co, ci, tap] (+)= sum_c partial[c, tap, ci, co], fixed order, four independent FP32 chains
__global__ void conv_wgrad_reduce_kernel(const float* __restrict__ partial, int nparts, int accumulate, float* __restrict__ gw) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;          // = (tap * 64 + ci) * 64 + co
    if (i >= kTaps * kC * kC) return;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    const size_t stride = (size_t)kTaps * kC * kC;
    int c = 0;
    for (; c + 4 <= nparts; c += 4) {
        s0 += __ldg(partial + (size_t)c * stride + i);
        s1 += __ldg(partial + (size_t)(c + 1) * stride + i);
        s2 += __ldg(partial + (size_t)(c + 2) * stride + i);
        s3 += __ldg(partial + (size_t)(c + 3) * stride + i);
    }
    for (; c < nparts; ++c) s0 += __ldg(partial + (size_t)c * stride + i);
    const float s = (s0 + s1) + (s2 + s3);
    const int co = i % kC, ci = (i / kC) % kC, tap = i / (kC * kC);
    float* dst = gw + ((size_t)co * kC + ci) * kTaps + tap;
    *dst = accumulate ? *dst + s : s;
}

int wgrad_plan(int N, int H, int W, int* tiles_per_cta, long long* grid) {
    const long long HW = (long long)H * W;
    const long long tpi = ceil_div(HW, kTileM), total = (long long)N * tpi;
    // T tiles per CTA.  Bound: <= 672 truncating MMAs per accumulator (12 per live sub-chunk), i.e.
    // T * (live sub-chunks per tile, averaged over an image) <= 56 -- 14 tiles of a 28x28 image, 16 of
    // a 14x14 one.  One CTA per SM is resident, so the launch costs waves * T tile-times: pick the T
    // that minimises it (14x14, N = 3200: T = 14 needs 458 CTAs = 4 waves, T = 15 needs 427 = 3).
    const long long sub_per_image = ceil_div(HW, 32);
    long long tmax = 56 * tpi / sub_per_image;
    if (tmax > 16) tmax = 16;
    if (tmax < 1) tmax = 1;
    long long T = tmax, best = -1;
    if (ceil_div(total, tmax) < kNumSMs) {
        T = ceil_div(total, kNumSMs);                   // small problem: spread over the SMs
    } else {
        for (long long t = tmax; t >= (tmax + 1) / 2; --t) {
            // + 1: a CTA's fixed cost (prologue, 147 KB partial written and re-read by the reduction)
            const long long cost = ceil_div(ceil_div(total, t), kNumSMs) * (t + 1);
            if (best < 0 || cost < best) { best = cost; T = t; }
        }
    }
    if (T < 1) T = 1;
    *tiles_per_cta = (int)T;
    *grid = ceil_div(total, T);                         // number of chunks
    return 0;
}

}  // namespace

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_conv3x3_c64_wgrad_usable(int N, int H, int W) {
    if (N <= 0 || H <= 0 || W <= 0) return 0;
    const long long HW = (long long)H * W;
    const int need = kTileM + ((W + 4) & ~3) + (W + 1);
    return (HW % 4 == 0) && need <= 228 && (long long)N * kC < (1LL << 31) && HW < (1LL << 24);
}

extern "C" size_t cmf_conv3x3_c64_wgrad_workspace(int N, int H, int W) {
    if (N <= 0 || H <= 0 || W <= 0) return 0;
    int T; long long chunks;
    wgrad_plan(N, H, W, &T, &chunks);
    const long long grid = chunks < kNumSMs ? chunks : kNumSMs;      // persistent: one partial per CTA
    return (size_t)grid * kTaps * kC * kC * sizeof(float);
}

extern "C" int cmf_conv3x3_c64_wgrad(int N, int H, int W, const float* x, const float* g_y, float* workspace,
                                     int accumulate, float* g_w, cmf_stream_t stream) {
    return cmf_conv3x3_c64_wgrad_det(N, H, W, x, g_y, workspace, accumulate, g_w, 0, stream);
}

extern "C" int cmf_conv3x3_c64_wgrad_det(int N, int H, int W, const float* x, const float* g_y, float* workspace,
                                         int accumulate, float* g_w, int deterministic, cmf_stream_t stream) {
    CMF_REQUIRE(N >= 0 && H >= 0 && W >= 0, "conv3x3_wgrad: negative size");
    CMF_REQUIRE(g_w, "conv3x3_wgrad: null g_w");
    if (N == 0 || H == 0 || W == 0) {
        if (!accumulate) {
            cudaError_t e = cudaMemsetAsync(g_w, 0, sizeof(float) * kTaps * kC * kC, (cudaStream_t)stream);
            if (e != cudaSuccess) { set_error("conv3x3_wgrad memset: %s", cudaGetErrorString(e)); return (int)e; }
        }
        return 0;
    }
    CMF_REQUIRE(x && g_y && workspace, "conv3x3_wgrad: null pointer");
    CMF_REQUIRE(cmf_conv3x3_c64_wgrad_usable(N, H, W), "conv3x3_wgrad: unsupported image size %dx%d (need H*W %% 4 == 0, W <= 47)", H, W);
    CMF_REQUIRE(aligned16(x) && aligned16(g_y) && aligned16(workspace), "conv3x3_wgrad: x, g_y and workspace must be 16-byte aligned");
    WgradParams P{};
    P.N = N; P.H = H; P.W = W; P.HW = H * W;
    P.tiles_per_image = (int)ceil_div(P.HW, kTileM);
    P.total_tiles = (long long)N * P.tiles_per_image;
    wgrad_plan(N, H, W, &P.tiles_per_cta, &P.num_chunks);
    const long long grid = P.num_chunks < kNumSMs ? P.num_chunks : kNumSMs;
    P.hl = (W + 4) & ~3;
    P.partial = workspace;
    const int pitch = (kTileM + P.hl + W + 1 <= 196) ? 196 : 228;
    CUtensorMap tmx, tmg;
    int rc = make_tmap_2d_plain_f32(&tmx, x, (uint64_t)N * kC, (uint64_t)P.HW, (uint64_t)P.HW, (uint32_t)pitch, kC);
    if (rc != 0) return rc;
    rc = make_tmap_2d_f32(&tmg, g_y, (uint64_t)N * kC, (uint64_t)P.HW, (uint64_t)P.HW, kC, 0);
    if (rc != 0) return rc;
    const size_t smem = (size_t)3 * kSub * kBBytes + (size_t)2 * kC * pitch * 4 + 1024;
    static DeviceOnce attr_set;
    if (!attr_set.done()) {
        cudaError_t e = cudaSuccess;
        const void* fns[4] = {(const void*)conv3x3_wgrad_tc_kernel<196, false>, (const void*)conv3x3_wgrad_tc_kernel<228, false>,
                              (const void*)conv3x3_wgrad_tc_kernel<196, true>, (const void*)conv3x3_wgrad_tc_kernel<228, true>};
        for (int f = 0; f < 4 && e == cudaSuccess; ++f)
            e = cudaFuncSetAttribute(fns[f], cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        if (e != cudaSuccess) { set_error("conv3x3_wgrad attr: %s", cudaGetErrorString(e)); return (int)e; }
        attr_set.mark();
    }
    const dim3 g((unsigned)grid), b(kWgThreads);
    cudaStream_t st = (cudaStream_t)stream;
    if (pitch == 196 && !deterministic) conv3x3_wgrad_tc_kernel<196, false><<<g, b, smem, st>>>(tmx, tmg, P);
    else if (pitch == 196) conv3x3_wgrad_tc_kernel<196, true><<<g, b, smem, st>>>(tmx, tmg, P);
    else if (!deterministic) conv3x3_wgrad_tc_kernel<228, false><<<g, b, smem, st>>>(tmx, tmg, P);
    else conv3x3_wgrad_tc_kernel<228, true><<<g, b, smem, st>>>(tmx, tmg, P);
    CMF_CHECK_LAUNCH();
    conv_wgrad_reduce_kernel<<<(kTaps * kC * kC + 255) / 256, 256, 0, (cudaStream_t)stream>>>(workspace, (int)grid, accumulate, g_w);
    CMF_CHECK_LAUNCH();
    return 0;
}
