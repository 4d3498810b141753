// 3x3 convolution, 64 -> 64 channels, stride 1, zero padding 1, NCHW fp32, on tcgen05 with 3xTF32.
//
// This is the conditioner hot spot of the image configs: every residual block of
// the coupler ResNet (`networks.py:35-94,116-161` of the reference) applies two such convolutions to
// the primal AND to all d tangent columns (`jvp_layers.py:49-64`), i.e. to B*(1+d) images per call.
//
// Formulation (implicit GEMM, nothing is materialised):
//     y[n, co, p] = sum_{tap, ci} x[n, ci, p + dy*W + dx] * w[co, ci, tap]      (p = flat pixel)
//   M = 128 consecutive pixels of the PACKED pixel stream (image 0's H*W pixels, then image 1's,
//   ...: a tile may straddle images, so no MMA row is wasted on the ragged end of an image --
//   28x28 needs 6.125 tiles per image instead of 7, 14x14 1.53 instead of 2), N = 64 output
//   channels, K = 9 taps x 64 inputs.
//   * The pixel operand lives in TENSOR MEMORY.  Per (tile, 32-channel chunk) the producer warp
//     (lane = channel) copies the tile's window of the stream -- the tile plus W+1 pixels either
//     side -- into shared memory as [channel][pixel] with 16-byte cp.async copies (a TMA box cannot
//     put two images' pixels in one row).  A "row worker" thread owns one output pixel: for each of the 9
//     taps it reads its 32 shifted inputs (or zero when the tap leaves ITS image: y + dy outside
//     [0, H) or x + dx outside [0, W) -- so whatever the window holds beyond an image border is
//     never used), splits them exactly into a TF32 head and a remainder (v = hi + lo) and writes
//     both to TMEM with tcgen05.st.  The MMA therefore reads only the weights from shared memory
//     (2 KB per instruction instead of 6 KB): the tensor pipe is no longer bound by shared-memory
//     bandwidth as the SS-mode Gram is.
//   * The weights arrive pre-split (hi | lo, tap-major, K-major rows) from a tiny prep kernel and
//     are streamed by TMA through a 4-deep ring; each stage is shared by the two pixel tiles of
//     the CTA, which halves the L2 -> SM traffic (the L2 slice throughput, not HBM, is the limit
//     once the pixel operand is read once instead of 9 times).
//   * Three MMAs per K-step: hi*Whi + hi*Wlo + lo*Whi (lo*Wlo ~ 2^-22 relative is dropped, as in
//     the Gram kernel).  The tensor core's FP32 accumulator truncates, so the accumulator is
//     drained into registers with round-to-nearest adds after every stage (K = 32).
//   * The same kernel is the data gradient: dgrad = conv(g_y, w^T flipped); the prep kernel writes
//     that variant of the weights.
#include "common.cuh"
#include "tc_common.cuh"
#include <cstdlib>

namespace cmf {
using namespace tc;

namespace {

// timing-only ablations (WRONG results): 1 = no tcgen05.ld in the drain, 2 = no tcgen05.st staging,
// 4 = no shared-memory loads / split, 8 = no MMAs (commit only), 16 = no epilogue stores
#ifndef CMF_CONV_ABLATE
#define CMF_CONV_ABLATE 0
#endif
// experiment switches: pure-spin (test_wait) instead of suspending try_wait on the critical hops
#ifndef CMF_CONV_SPIN
#define CMF_CONV_SPIN 0
#endif
// Experiment switch (measured, rejected): a converged warp waits on an mbarrier through ONE lane and is
// released by __syncwarp, to cut the number of mbarrier transactions (16 worker warps x 32 lanes poll
// two `done` phases per stage).  It is SLOWER (1.149 vs 1.056 ms on the 28x28 tangent launch): the
// polls are not what the synchronisation skeleton costs (profiles/r02_summary.md).
#ifndef CMF_CONV_WARPWAIT
#define CMF_CONV_WARPWAIT 0
#endif
__device__ __forceinline__ void wait_warp(uint64_t* bar, uint32_t parity) {
    if (CMF_CONV_WARPWAIT) {
        if ((threadIdx.x & 31) == 0) mbar_wait(bar, parity);
        __syncwarp();
    } else {
        mbar_wait(bar, parity);
    }
}
__device__ __forceinline__ void wait_issuer(uint64_t* bar, uint32_t parity) {
    if (CMF_CONV_SPIN & 1) mbar_spin(bar, parity); else wait_warp(bar, parity);
}
__device__ __forceinline__ void wait_worker(uint64_t* bar, uint32_t parity) {
    if (CMF_CONV_SPIN & 2) mbar_spin(bar, parity); else wait_warp(bar, parity);
}

constexpr int kC = 64;                 // channels (in = out)
constexpr int kTileM = 128;            // output pixels per tile (TMEM lanes)
constexpr int kTiles = 2;              // tiles per CTA sharing one weight stream
constexpr int kChunkC = 32;            // input channels per stage (K = 32)
constexpr int kTaps = 9;
constexpr int kStages = 2 * kTaps;     // per tile: 2 channel chunks x 9 taps
constexpr int kWRing = 4;
constexpr int kWStageBytes = 2 * kC * kChunkC * 4;     // hi + lo tile: 16 KB
// Row workers: kGroups warps per TMEM lane quarter and tile.  Group g of a tile stages channels
// [16 g, 16 g + 16) of every 32-channel stage and drains / stores output channels [32 g, 32 g + 32): the
// work per stage is unchanged, but each scheduler now holds four worker warps instead of two, which is
// what hides the tcgen05.st / tcgen05.ld / mbarrier latencies the workers are bound by (r01: 0.8
// eligible warps per scheduler, tensor pipe 46 % active).
#ifndef CMF_CONV_GROUPS
#define CMF_CONV_GROUPS 2
#endif
constexpr int kGroups = CMF_CONV_GROUPS;
constexpr int kWorkerWarps = 4 * kGroups * kTiles;      // 16
constexpr int kMmaWarp1 = 2 + kWorkerWarps;             // second MMA issuer (tile 1)
constexpr int kHaloWarp = 3 + kWorkerWarps;              // halo-window loader
constexpr int kConvThreads = 32 * (4 + kWorkerWarps);   // 640
constexpr int kCg = kChunkC / kGroups;                  // channels of a stage staged by one worker
constexpr int kOg = kC / kGroups;                       // output channels drained by one worker
constexpr int kConvTmemCols = 512;
constexpr int kMaxBox = 256;

__device__ __forceinline__ float ld_shared_f32(uint32_t addr) {
    float v;
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

#ifdef CMF_CONV_PROF
__device__ long long g_conv_prof[128];
#define PROF(i) do { if (blockIdx.x == 70 && lane == 0 && (i) < 128) g_conv_prof[(i)] = clock64(); } while (0)
#else
#define PROF(i) do { } while (0)
#endif

struct ConvParams {
    int N, H, W, HW;
    long long total_px;                // N * H * W: length of the packed pixel stream
    long long total_tiles;             // ceil(total_px / 128)
    int hl;                            // halo pixels in front of the tile: W+1 rounded up to 4 (bulk copies need 16-byte starts)
    int win;                           // window length in pixels: hl + 128 + W + 1 rounded up to 4 (<= kPitch)
    const float* x;                    // (N, 64, H, W)
    const float* bias;                 // (64) or null
    // fused epilogue:  t = acc + bias;  if (mask_first) t *= m;  if (res) t += res;  y = t;
    //                  if (y2) y2 = t * m      with m = mask_src[n / mask_div, co, p] > 0
    const float* mask_src;             // (N / mask_div, 64, H, W) pre-ReLU primal activations, or null
    const uint2* mask_bits;            // (N / mask_div, H*W) packed signs of the same (cmf_relu_mask_pack), or null
    int mask_div;
    int tail;                          // joint batch: images [tail, N) are the PRIMAL images of samples 0 .. N - tail - 1
                                       // (mask row im - tail, and the only images that get the bias); 0 = no tail
    int mask_first;
    const float* res;                  // (N, 64, H, W) or null
    float* y;
    float* y2;                         // (N, 64, H, W) or null
};

// weight prep: w (co, ci, 3, 3) -> out[part][tap][row][col] with part 0 = TF32 head, 1 = remainder.
// transpose == 0: row = co, col = ci, tap = ky*3+kx            (forward / tangent push)
// transpose == 1: row = ci, col = co, tap = (2-ky)*3 + (2-kx)  (data gradient)
__global__ void conv_prep_kernel(const float* __restrict__ w, int transpose, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= kTaps * kC * kC) return;
    const int col = i % kC, row = (i / kC) % kC, tap = i / (kC * kC);
    int co, ci, t;
    if (!transpose) { co = row; ci = col; t = tap; }
    else { ci = row; co = col; t = kTaps - 1 - tap; }
    const float v = __ldg(w + ((long long)co * kC + ci) * kTaps + t);
    // round-to-nearest head: the remainder is signed and half as large, so the tensor core's
    // truncation of it to TF32 is unbiased
    const float hi = __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xFFFFE000u);
    const float lo = v - hi;
    out[i] = hi;
    out[kTaps * kC * kC + i] = lo;
}

// bits[n, p] = {bit e: x[n, e, p] > 0 (channels 0..31), bit e: x[n, 32 + e, p] > 0}: the relu-JVP mask of
// a primal activation, packed once and reused by the d tangent images of the sample in the forward
// and the backward launches (64 strided loads per output pixel become one 8-byte load)
__global__ void relu_mask_pack_kernel(long long total, int hw, const float* __restrict__ x, uint2* __restrict__ bits) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;      // = n * hw + p
    if (i >= total) return;
    const long long n = i / hw;
    const float* src = x + (n * kC) * hw + (i - n * hw);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int e = 0; e < 32; ++e) {
        lo |= (__ldg(src + (long long)e * hw) > 0.f ? 1u : 0u) << e;
        hi |= (__ldg(src + (long long)(32 + e) * hw) > 0.f ? 1u : 0u) << e;
    }
    bits[i] = make_uint2(lo, hi);
}

// kPitch = pixels per halo row in shared memory = TMA box width (compile-time so that the 32
// channel loads of a row worker are LDS with immediate offsets): 192 covers W <= 31, 256 W <= 61.
//
// PERSISTENT: one CTA per SM walks tile pairs blockIdx.x, blockIdx.x + gridDim.x, ...  All mbarrier
// phases are functions of the global stage counter gs = 18 * it + s, so the weight ring, the
// staging buffers and the ping-pong accumulators simply keep rolling across pairs; the halo tile of
// the next pair is fetched while the current one is still being consumed (chunk 0 after stage 9 has
// completed, chunk 1 after stage 0 of the next pair).  Launch, TMEM allocation, barrier set-up and
// the first TMA round trip (a third of a one-pair CTA's life) are paid once per SM.
// kSeg = stages (K = 32 each) accumulated on the tensor core between two drains of the accumulator
// into the FP32 registers; kStages % kSeg == 0, so segments never straddle tile pairs.
// kFloatMask: the epilogue may be given the mask as float activations (cmf_conv3x3_c64_fused); the
// step itself only uses packed sign bits, and keeping the 64-load float path out of that
// instantiation keeps its code small (the kernel is sensitive to instruction-cache footprint:
// +20 KB of specialised epilogues once cost 10 % of every launch).
template <int kPitch, int kSeg, bool kFloatMask>
__global__ void __launch_bounds__(kConvThreads, 1)
conv3x3_tc_kernel(const __grid_constant__ CUtensorMap tmw, ConvParams P) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ uint64_t w_full[kWRing], done[kWRing];     // done[gs % 4]: the MMAs of stage gs have completed for both tiles
    __shared__ uint64_t a_full[kTiles][2];
    __shared__ uint64_t at_full[kTiles][2];
    __shared__ uint64_t halo_go;                           // producer -> halo warp: the next halo buffers are free
    __shared__ uint32_t tmem_base_smem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* wring = smem;                                         // kWRing x 16 KB, 1024-aligned
    float* halo = reinterpret_cast<float*>(smem + kWRing * kWStageBytes);   // [tile][chunk][32][kPitch]
    constexpr int kHaloElems = kChunkC * kPitch;

    const long long npairs = (P.total_tiles + kTiles - 1) / kTiles;
    const int nit = (int)((npairs - blockIdx.x + gridDim.x - 1) / gridDim.x);     // pairs of this CTA (>= 1)
    // tile t of this CTA's it-th pair: first pixel g0 of the packed stream
    auto tile_of = [&](int it, int t, bool& act, long long& g0) {
        const long long T = ((long long)blockIdx.x + (long long)it * gridDim.x) * kTiles + t;
        act = T < P.total_tiles;
        g0 = act ? T * kTileM : 0;
    };

    if (threadIdx.x == 0) {
        for (int r = 0; r < kWRing; ++r) { mbar_init(&w_full[r], 1); mbar_init(&done[r], kTiles); }
        mbar_init(&halo_go, 1);
        for (int t = 0; t < kTiles; ++t) {
            for (int c = 0; c < 2; ++c) {
                mbar_init(&a_full[t][c], 32);              // one cp.async arrival per producer lane
                mbar_init(&at_full[t][c], kGroups * kTileM / 32);     // one arrival per worker warp
            }
        }
        mbar_fence_init();
        prefetch_tmap(&tmw);
    }
    if (warp == 1) tmem_alloc<kConvTmemCols>(&tmem_base_smem);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_smem;
    // TMEM columns: [0,256) accumulators, (tile, stage parity) -> 64 cols; [256,512) operand staging,
    // (tile, buffer) -> 64 cols (hi | lo)
    constexpr uint32_t kAccCol = 0, kStgCol = 256;

    if (warp == 0) {
        // ------------------------------------------------------------------ weight producer (TMA)
        if (lane == 0) {
            int gs = 0;
            for (int it = 0; it < nit; ++it) {
                for (int s = 0; s < kStages; ++s, ++gs) {
                    const int r = gs % kWRing;
                    if (gs >= kWRing) mbar_wait(&done[r], ((gs / kWRing) - 1) & 1);      // stage gs - 4 has completed
                    const int chunk = s / kTaps, tap = s % kTaps;
                    uint8_t* dst = wring + (size_t)r * kWStageBytes;
                    if (CMF_CONV_ABLATE & 32) {
                        mbar_arrive(&w_full[r]);
                    } else {
                        mbar_arrive_expect_tx(&w_full[r], kWStageBytes);
                        tma_load_2d(dst, &tmw, &w_full[r], chunk * kChunkC, tap * kC);                       // hi
                        tma_load_2d(dst + kWStageBytes / 2, &tmw, &w_full[r], chunk * kChunkC, (kTaps + tap) * kC);   // lo
                    }
                    // halo buffers are recycled as soon as every worker is past their last read:
                    // chunk 1 of this pair once stage 0 of this pair has completed (s = 4 waited for it),
                    // chunk 0 of the next pair once stage 9 has completed (s = 13 waited for it).  The
                    // halo warp is told; it cannot fall more than one phase behind, because the next
                    // signal needs a stage that consumes the data of this one.
                    if (s == 4 || (s == 13 && it + 1 < nit)) mbar_arrive(&halo_go);
                }
            }
        }
    } else if (warp == kHaloWarp) {
        // ------------------------------------------------------------------ halo-window loader
        // All 32 lanes fetch the windows with 16-byte cp.async copies (LDGSTS).  Column j of a window
        // row is pixel g0 - hl + j of the packed stream, which is contiguous in memory only inside
        // one image: a lane owns the 4-pixel columns {lane, lane + 32} of the window, resolves their
        // (image, pixel) once and walks the 32 channel rows of the chunk -- consecutive lanes copy
        // consecutive 16 bytes of a row.  (One TMA box cannot place two images' pixels in one row;
        // one-row bulk copies cost ~0.25 ms per launch, and issuing the copies from the weight
        // producer holds up the weight stream: both measured.)  Completion is signalled per lane
        // with cp.async.mbarrier.arrive.noinc (a_full counts 32).
        auto load_halo = [&](int it, int c, int t) {
            bool act; long long g0;
            tile_of(it, t, act, g0);
            if (!act) return;
            const long long ws = g0 - P.hl;
            const uint32_t dst_buf = smem_u32(halo + (size_t)(t * 2 + c) * kHaloElems);
            const long long hw = P.HW;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int col = 4 * (lane + 32 * h);
                const long long sp = ws + col;
                if (!(CMF_CONV_ABLATE & 64) && col < P.win && sp >= 0 && sp < P.total_px) {
                    const long long im = sp / hw;
                    const float* src = P.x + ((im * kC + c * kChunkC) * hw + (sp - im * hw));
                    const uint32_t dst = dst_buf + (uint32_t)(col * 4);
#pragma unroll 8
                    for (int r = 0; r < kChunkC; ++r)
                        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (uint32_t)(r * kPitch * 4)),
                                     "l"(src + (long long)r * hw)
                                     : "memory");
                }
            }
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&a_full[t][c])) : "memory");
        };
        // load s of this CTA: chunk (s & 1) of pair (s >> 1); load 0 needs no permission, load s >= 1
        // waits for the producer's (s - 1)-th signal.  One call site (code footprint).
        for (int s = 0; s < 2 * nit; ++s) {
            if (s >= 1) mbar_wait(&halo_go, (s - 1) & 1);
#pragma unroll 1
            for (int t = 0; t < kTiles; ++t) load_halo(s >> 1, s & 1, t);
        }
    } else if (warp == 1 || warp == kMmaWarp1) {
        // ------------------------------------------------------------------ MMA issuers, one warp per tile
        // (the issue path of a single thread is latency-bound at ~50 cycles per tcgen05.mma; two
        // issuers keep the tensor pipe fed).  The whole warp walks the loop converged; one elected
        // lane issues.
        const int t = warp == 1 ? 0 : 1;
        constexpr uint32_t idesc = umma_idesc_tf32(kTileM, kC);
        const uint32_t acc0 = tmem_base + kAccCol + (uint32_t)(t * 2 * kC);
        const uint32_t wring_addr = smem_u32(wring);
        int gs = 0;
        for (int it = 0; it < nit; ++it) {
            bool act; long long g0;
            tile_of(it, t, act, g0);
            for (int s = 0; s < kStages; ++s, ++gs) {
                const int r = gs % kWRing;
                const int b = gs & 1;
                wait_issuer(&w_full[r], (gs / kWRing) & 1);
                if (act) wait_issuer(&at_full[t][b], (gs >> 1) & 1);
                // accumulator (segment gs / kSeg, parity) was drained (two segments ago) by every worker warp BEFORE it arrived on
                // at_full for this stage (program order + tcgen05.wait::ld), so no extra barrier is needed
                tc_fence_after();
                const uint32_t acc = acc0 + (uint32_t)(((gs / kSeg) & 1) * kC);
                const uint32_t fresh = (gs % kSeg) == 0 ? 0u : 1u;      // 0: first stage of a segment overwrites
                if (elect_one()) {
                    if (act && !(CMF_CONV_ABLATE & 8)) {
                        const uint64_t dbh = umma_desc_sw128(wring_addr + (uint32_t)(r * kWStageBytes));
                        const uint64_t dbl = umma_desc_sw128(wring_addr + (uint32_t)(r * kWStageBytes + kWStageBytes / 2));
                        const uint32_t a_hi = tmem_base + kStgCol + (uint32_t)((t * 2 + b) * 64);
                        const uint32_t a_lo = a_hi + 32;
                        // cross terms first: they land in a fresh (small) accumulator, where the tensor
                        // core's truncating FP32 add costs nothing; only the three hi*hi adds that follow
                        // see a large accumulator
#pragma unroll
                        for (int k = 0; k < kChunkC / 8; ++k) {      // +32 bytes along K = +2 in the descriptor's address field
                            umma_tf32_ts(acc, a_hi + k * 8, dbl + 2 * k, idesc, (k != 0) | fresh);
                            umma_tf32_ts(acc, a_lo + k * 8, dbh + 2 * k, idesc, 1);
                        }
#pragma unroll
                        for (int k = 0; k < kChunkC / 8; ++k) umma_tf32_ts(acc, a_hi + k * 8, dbh + 2 * k, idesc, 1);
                    }
                    if (CMF_CONV_ABLATE & 256) mbar_arrive(&done[r]);
                    else umma_commit(&done[r]);      // frees the weight slot and this tile's staging buffer; ends the segment
                }
                __syncwarp();
            }
        }
    } else {
        // ------------------------------------------------------------------ row workers
        const int wi = warp - 2;
        const int t = wi / (4 * kGroups);
        const int grp = (wi >> 2) % kGroups;           // channel group of this warp (see kGroups)
        const int rb = warp & 3;                       // TMEM lane quarter this warp may touch
        const int m = rb * 32 + lane;                  // row of the tile = output pixel p0 + m
        const uint32_t lane_addr = (uint32_t)(rb * 32) << 16;
        const uint32_t row_addr = smem_u32(halo) + (uint32_t)((t * 2 * kHaloElems + grp * kCg * kPitch + P.hl + m) * 4);
        const long long hw = P.HW;
        constexpr int kHalves = kOg / 32;              // 32-column accumulator slices this worker owns
        for (int it = 0; it < nit; ++it) {
            bool act; long long g0;
            tile_of(it, t, act, g0);
            if (!act) break;                           // only the very last pair can lack its second tile
            const int gs0 = it * kStages;
            // this worker's pixel of the packed stream -> (image, flat pixel, row, column)
            const long long g = g0 + m;
            const bool valid = g < P.total_px;         // the stream's last tile may be ragged
            const int im = valid ? (int)(g / hw) : 0;
            const int p = valid ? (int)(g - (long long)im * hw) : 0;
            const int yy = p / P.W, xx = p - yy * P.W;
            // bit tap of okbits: the tap stays inside this pixel's image (otherwise the operand is zero;
            // the window holds a neighbouring image's pixels or nothing at that position)
            uint32_t okbits = 0;
            if (valid) {
#pragma unroll
                for (int tp = 0; tp < kTaps; ++tp) {
                    const int dy = tp / 3 - 1, dx = tp % 3 - 1;
                    const bool ok = yy + dy >= 0 && yy + dy < P.H && xx + dx >= 0 && xx + dx < P.W;
                    okbits |= (ok ? 1u : 0u) << tp;
                }
            }
            float accr[kOg];
#pragma unroll
            for (int j = 0; j < kOg; ++j) accr[j] = 0.f;
            // software pipeline: the shared-memory loads of stage s+1 are issued while the
            // tcgen05.st of stage s is in flight (its completion wait is the longest latency here)
            float v[kCg];
            auto load_stage = [&](int s) {
                const int chunk = s / kTaps, tap = s - chunk * kTaps;
                const int dy = tap / 3 - 1, dx = tap - (tap / 3) * 3 - 1;
                if (tap == 0) wait_worker(&a_full[t][chunk], it & 1);
                const uint32_t src = row_addr + (uint32_t)((chunk * kHaloElems + dy * P.W + dx) * 4);
                const bool ok = (okbits >> tap) & 1u;
                if (CMF_CONV_ABLATE & 4) {
#pragma unroll
                    for (int c = 0; c < kCg; ++c) v[c] = (float)(c + tap);
                } else if (__all_sync(0xffffffffu, ok)) {      // interior rows: no zero-fill, no predicates
#pragma unroll
                    for (int c = 0; c < kCg; ++c) v[c] = ld_shared_f32(src + (uint32_t)(c * kPitch * 4));
                } else {
#pragma unroll
                    for (int c = 0; c < kCg; ++c) v[c] = ok ? ld_shared_f32(src + (uint32_t)(c * kPitch * 4)) : 0.f;
                }
            };
            // drain the accumulator of the segment that ended with global stage e into the FP32 registers
            auto drain = [&](int e) {
                wait_worker(&done[e % kWRing], (e / kWRing) & 1);
                tc_fence_after();
                const uint32_t acc = tmem_base + kAccCol + (uint32_t)((t * 2 + ((e / kSeg) & 1)) * kC + grp * kOg) + lane_addr;
#pragma unroll
                for (int h = 0; h < kHalves; ++h) {
                    float u[32];
                    if (CMF_CONV_ABLATE & 1) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) u[j] = accr[h * 32 + j] * 0.5f;
                    } else {
                        tmem_ld_32x32(acc + h * 32, u);
                    }
#pragma unroll
                    for (int j = 0; j < 32; j += 2) fadd2(accr[h * 32 + j], accr[h * 32 + j + 1], u[j], u[j + 1]);
                }
                tc_fence_before();
            };
            load_stage(0);
            // One extra trip: iteration kStages only drains the pair's last segment, so the drain has a
            // single call site (the kernel is sensitive to its instruction-cache footprint).
            for (int s = 0; s <= kStages; ++s) {
                const int gs = gs0 + s;
                if (s < kStages) {
                    const int b = gs & 1;
                    float hi[kCg], lo[kCg];
#pragma unroll
                    for (int c = 0; c < kCg; ++c) {
                        // TF32 head: round to nearest where the result feeds the metric (kSeg == 1); the
                        // gradient launches (kSeg >= 2) truncate -- one operation less, the remainder then
                        // loses 2 of its 13 bits in the tensor core (2^-21 |x| instead of 2^-22)
                        hi[c] = __uint_as_float((__float_as_uint(v[c]) + (kSeg == 1 ? 0x1000u : 0u)) & 0xFFFFE000u);
                        lo[c] = v[c] - hi[c];                                                       // exact signed remainder
                    }
                    if (gs >= 2) wait_worker(&done[(gs - 2) % kWRing], ((gs - 2) / kWRing) & 1);     // staging buffer b is free
                    tc_fence_after();
                    const uint32_t stg = tmem_base + kStgCol + (uint32_t)((t * 2 + b) * 64 + grp * kCg) + lane_addr;
                    if (!(CMF_CONV_ABLATE & 2)) {
                        tmem_st_cols<kCg>(stg, hi);
                        tmem_st_cols<kCg>(stg + 32, lo);
                    } else {
                        float keep = 0.f;
#pragma unroll
                        for (int c = 0; c < kCg; ++c) keep += hi[c] + lo[c];
                        if (keep == 123.456f) accr[0] += keep;      // keeps the split alive
                    }
                    if (s + 1 < kStages) load_stage(s + 1);
                    tmem_st_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&at_full[t][b]);
                }
                // drain one stage late, so the tensor pipe already works on the next stage in the other
                // accumulator; the pair's last segment right after its own stage
                if (!(CMF_CONV_ABLATE & 128) && s > 0 && ((gs % kSeg) == 0 || s == kStages)) drain(gs - 1);
            }
            if ((CMF_CONV_ABLATE & 16) && valid) {
                float keep = 0.f;
#pragma unroll
                for (int j = 0; j < kOg; ++j) keep += accr[j];
                if (keep == 123.456f) P.y[0] = keep;
            } else if (valid) {
                const int j0 = grp * kOg;                                  // first output channel of this worker
                const long long o = (long long)im * kC * hw + p;
                // joint batch: tangent images first (mask row im / mask_div, no bias), then the primal images
                const bool in_tail = P.tail > 0 && im >= P.tail;
                const int mrow = in_tail ? im - P.tail : im / P.mask_div;
                const float* bias = (P.tail == 0 || in_tail) ? P.bias : nullptr;
                if (!P.mask_src && !P.mask_bits && !P.res) {
#pragma unroll
                    for (int j = 0; j < kOg; ++j)
                        P.y[o + (long long)(j0 + j) * hw] = accr[j] + (bias ? __ldg(bias + j0 + j) : 0.f);
                } else {
                    const float* ms = P.mask_src ? P.mask_src + (long long)mrow * kC * hw + p : nullptr;
                    const float* rs = P.res ? P.res + o : nullptr;
                    float* y1 = P.y + o;
                    float* y2 = P.y2 ? P.y2 + o : nullptr;
                    const bool mask_first = P.mask_first != 0;
                    // every load of a batch is issued before the batch's stores so that the latencies
                    // overlap (the compiler cannot hoist loads over possibly aliasing stores): first the
                    // mask values, packed into words, then the skip connection 32 channels at a time
                    uint32_t keep_bits[2] = {0xffffffffu, 0xffffffffu};
                    if (P.mask_bits) {       // one 8-byte load instead of 64 strided ones
                        const uint2 mb = __ldg(P.mask_bits + (long long)mrow * hw + p);
                        keep_bits[0] = mb.x;
                        keep_bits[1] = mb.y;
                    } else if (kFloatMask && ms) {
#pragma unroll
                        for (int h = 0; h < kHalves; ++h) {
                            const int hh = j0 / 32 + h;
                            float mv[32];
#pragma unroll
                            for (int e = 0; e < 32; ++e) mv[e] = __ldg(ms + (32 * hh + e) * hw);
                            uint32_t bits = 0;
#pragma unroll
                            for (int e = 0; e < 32; ++e) bits |= (mv[e] > 0.f ? 1u : 0u) << e;
                            keep_bits[hh] = bits;
                        }
                    }
#pragma unroll
                    for (int h = 0; h < kHalves; ++h) {
                        const int hh = j0 / 32 + h;                     // 32-channel word of the mask / slice of the output
                        float rv[32];
#pragma unroll
                        for (int e = 0; e < 32; ++e) rv[e] = rs ? __ldg(rs + (32 * hh + e) * hw) : 0.f;
#pragma unroll
                        for (int e = 0; e < 32; ++e) {
                            const int j = 32 * hh + e;
                            float v2 = accr[32 * h + e] + (bias ? __ldg(bias + j) : 0.f);
                            const bool keep = (keep_bits[hh] >> e) & 1u;
                            if (mask_first && !keep) v2 = 0.f;
                            v2 += rv[e];
                            y1[j * hw] = v2;
                            if (y2) y2[j * hw] = keep ? v2 : 0.f;
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
        tmem_dealloc<kConvTmemCols>(tmem_base);
    }
}

}  // namespace

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_conv3x3_c64_prep(const float* w, int transpose, float* wsplit, cmf_stream_t stream) {
    CMF_REQUIRE(w && wsplit, "conv_prep: null pointer");
    const int n = kTaps * kC * kC;
    conv_prep_kernel<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(w, transpose != 0, wsplit);
    CMF_CHECK_LAUNCH();
    return 0;
}

extern "C" int cmf_conv3x3_c64_usable(int N, int H, int W) {
    if (N <= 0 || H <= 0 || W <= 0) return 0;
    const long long HW = (long long)H * W;
    const int need = kTileM + ((W + 4) & ~3) + (W + 1);        // halo box: front halo + tile + back halo
    return (HW % 4 == 0) && need <= kMaxBox && (long long)N * kC < (1LL << 31) && HW < (1LL << 24);
}

static int conv_launch(int N, int H, int W, const float* x, const float* wsplit, const float* bias,
                       const float* mask_src, const uint32_t* mask_bits, int mask_div, int mask_first, const float* res,
                       float* y, float* y2, int drain_stages, cmf_stream_t stream, int tail = 0) {
    CMF_REQUIRE(drain_stages >= 1 && drain_stages <= 3, "conv3x3: drain_stages must be 1, 2 or 3 (got %d)", drain_stages);
    CMF_REQUIRE(N >= 0 && H >= 0 && W >= 0, "conv3x3: negative size");
    if (N == 0 || H == 0 || W == 0) return 0;
    CMF_REQUIRE(x && wsplit && y, "conv3x3: null pointer");
    CMF_REQUIRE(cmf_conv3x3_c64_usable(N, H, W), "conv3x3: unsupported image size %dx%d (need H*W %% 4 == 0, W <= 61)", H, W);
    CMF_REQUIRE(aligned16(x) && aligned16(wsplit), "conv3x3: x and wsplit must be 16-byte aligned");
    ConvParams P{};
    P.N = N; P.H = H; P.W = W; P.HW = H * W;
    P.total_px = (long long)N * P.HW;
    P.total_tiles = ceil_div(P.total_px, (long long)kTileM);
    P.x = x;
    P.hl = (W + 4) & ~3;
    const bool masked = mask_src || mask_bits;
    CMF_REQUIRE(tail >= 0 && tail <= N, "conv3x3: tail must lie in [0, N]");
    CMF_REQUIRE(!masked || (mask_div >= 1 && (tail > 0 ? tail : N) % mask_div == 0), "conv3x3: mask_div must divide the tangent images");
    CMF_REQUIRE(tail == 0 || !masked || tail / mask_div == N - tail, "conv3x3: joint batch needs tail / mask_div == N - tail primal images");
    P.tail = tail;
    CMF_REQUIRE(!y2 || masked, "conv3x3: second output needs a mask source");
    CMF_REQUIRE(!mask_bits || (reinterpret_cast<uintptr_t>(mask_bits) & 7u) == 0, "conv3x3: mask_bits must be 8-byte aligned");
    P.bias = bias; P.y = y; P.y2 = y2;
    P.mask_src = mask_bits ? nullptr : mask_src; P.mask_bits = reinterpret_cast<const uint2*>(mask_bits);
    P.mask_div = masked ? mask_div : 1; P.mask_first = mask_first; P.res = res;
    P.win = (kTileM + P.hl + W + 1 + 3) & ~3;
    const int pitch = P.win <= 192 ? 192 : 256;
    CUtensorMap tmw;
    int rc = make_tmap_2d_f32(&tmw, wsplit, (uint64_t)2 * kTaps * kC, (uint64_t)kC, (uint64_t)kC, kC, 0);
    if (rc != 0) return rc;
    const size_t smem = (size_t)kWRing * kWStageBytes + (size_t)kTiles * 2 * kChunkC * pitch * 4 + 1024;
    // persistent: one CTA per SM walks the tile pairs
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long npairs = ceil_div(P.total_tiles, kTiles);
    const unsigned grid = (unsigned)(npairs < sms ? npairs : sms);
    // CMF_CONV_SEG (debug): force one drain interval for every launch
    static int seg_env = -1;
    if (seg_env < 0) { const char* e = getenv("CMF_CONV_SEG"); seg_env = e ? atoi(e) : 0; if (seg_env < 0 || seg_env > 3) seg_env = 0; }
    const int seg = seg_env ? seg_env : drain_stages;
#define CMF_CONV_LAUNCH(PITCH, SEG, FM)                                                                                   \
    do {                                                                                                               \
        static DeviceOnce attr_set;                                                                                  \
        if (!attr_set.done()) {                                                                                               \
            cudaError_t e = cudaFuncSetAttribute(conv3x3_tc_kernel<PITCH, SEG, FM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); \
            if (e != cudaSuccess) { set_error("conv3x3 attr: %s", cudaGetErrorString(e)); return (int)e; }           \
            attr_set.mark();                                                                                           \
        }                                                                                                              \
        conv3x3_tc_kernel<PITCH, SEG, FM><<<grid, kConvThreads, smem, (cudaStream_t)stream>>>(tmw, P);                 \
    } while (0)
    if (P.mask_src) {          // float mask: rare (tests, external callers); per-stage drain only
        CMF_REQUIRE(seg == 1, "conv3x3: a float mask_src needs drain_stages == 1");
        if (pitch == 192) CMF_CONV_LAUNCH(192, 1, true); else CMF_CONV_LAUNCH(256, 1, true);
    } else if (pitch == 192) { if (seg == 1) CMF_CONV_LAUNCH(192, 1, false); else if (seg == 2) CMF_CONV_LAUNCH(192, 2, false); else CMF_CONV_LAUNCH(192, 3, false); }
    else { if (seg == 1) CMF_CONV_LAUNCH(256, 1, false); else if (seg == 2) CMF_CONV_LAUNCH(256, 2, false); else CMF_CONV_LAUNCH(256, 3, false); }
#undef CMF_CONV_LAUNCH
    CMF_CHECK_LAUNCH();
    return 0;
}

#ifdef CMF_CONV_PROF
extern "C" int cmf_debug_conv_prof(long long* host_out) {
    return (int)cudaMemcpyFromSymbol(host_out, g_conv_prof, sizeof(long long) * 128);
}
#endif

extern "C" int cmf_conv3x3_c64_fused(int N, int H, int W, const float* x, const float* wsplit, const float* bias,
                                     const float* mask_src, int mask_div, int mask_first, const float* res,
                                     float* y, float* y2, cmf_stream_t stream) {
    return conv_launch(N, H, W, x, wsplit, bias, mask_src, nullptr, mask_div, mask_first, res, y, y2, 1, stream);
}

extern "C" int cmf_conv3x3_c64_fused_bits(int N, int H, int W, const float* x, const float* wsplit, const float* bias,
                                          const uint32_t* mask_bits, int mask_div, int mask_first, const float* res,
                                          float* y, float* y2, int drain_stages, cmf_stream_t stream) {
    return conv_launch(N, H, W, x, wsplit, bias, nullptr, mask_bits, mask_div, mask_first, res, y, y2, drain_stages, stream);
}

extern "C" int cmf_conv3x3_c64_joint(int N, int Nt, int H, int W, const float* x, const float* wsplit, const float* bias,
                                     const uint32_t* mask_bits, int mask_div, int mask_first, const float* res,
                                     float* y, float* y2, int drain_stages, cmf_stream_t stream) {
    CMF_REQUIRE(Nt > 0 && Nt < N, "conv3x3_joint: need 0 < Nt < N (Nt tangent images, then N - Nt primal images)");
    return conv_launch(N, H, W, x, wsplit, bias, nullptr, mask_bits, mask_div, mask_first, res, y, y2, drain_stages, stream, Nt);
}

extern "C" int cmf_relu_mask_pack(int N, int HW, const float* x, uint32_t* bits, cmf_stream_t stream) {
    CMF_REQUIRE(N >= 0 && HW >= 0, "relu_mask_pack: negative size");
    if (N == 0 || HW == 0) return 0;
    CMF_REQUIRE(x && bits, "relu_mask_pack: null pointer");
    CMF_REQUIRE((reinterpret_cast<uintptr_t>(bits) & 7u) == 0, "relu_mask_pack: bits must be 8-byte aligned");
    const long long total = (long long)N * HW;
    relu_mask_pack_kernel<<<(unsigned)ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(total, HW, x, reinterpret_cast<uint2*>(bits));
    CMF_CHECK_LAUNCH();
    return 0;
}

extern "C" int cmf_conv3x3_c64_fwd(int N, int H, int W, const float* x, const float* wsplit, const float* bias,
                                   float* y, cmf_stream_t stream) {
    return cmf_conv3x3_c64_fused(N, H, W, x, wsplit, bias, nullptr, 1, 0, nullptr, y, nullptr, stream);
}
