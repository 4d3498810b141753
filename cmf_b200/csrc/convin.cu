// First convolution of the image couplers' conditioner: C_in -> 64 channels, 3 x 3, padding 1, no bias
// (C_in = 1 .. 4: the passthrough half of a 1- to 8-channel image level), forward / tangent push, data
// gradient and weight gradient.
//
// Replaces cmf/models/components/networks.py:131-139 (the first `get_conv` of get_resnet) and its JVP
// rule jvp_layers.py:49-64 (the same bias-free convolution applied to every tangent image), which the
// reference and rounds 1-2 of this build ran on cuDNN (implicit_convolve_sgemm, dgrad_engine,
// wgrad_alg0_engine + FFT variants).
//
// With K = 9 C_in <= 36 this is not a tensor-core shape: all three kernels are bound by the 64-channel
// tensor they write or read (256 B per pixel), and are laid out so that this tensor is touched once with
// coalesced accesses:
//   forward        a thread owns an output pixel and its 64 accumulators; the 9 C_in taps come through L1;
//   data gradient  two phases per image band: z[k][p] = sum_co w[co][k] g_y[co][p] (each g_y value read once),
//                  then g_x[ci][q] = sum_tap z[ci, tap][q - shift(tap)] out of shared memory;
//   weight grad    per band, g_y transposed in shared memory; a warp walks pixels, lane = (8 output
//                  channels) x (tap row), 24 C_in accumulators per lane; persistent blocks, one partial per
//                  block, reduced in block order (bitwise reproducible).
#include "common.cuh"

namespace cmf {
namespace {

constexpr int kCo = 64;
constexpr int kThreads = 256;
constexpr int kMaxCin = CMF_CONVIN_MAX_CIN;     // 16
constexpr int kChunk = 4;                       // input channels per pass of the gradient kernels

struct InParams {
    int N, Cin, H, W, R, nb;       // R rows per band, nb bands per image
    int CinT, ci0;                 // channels of the whole tensor / first channel of this pass (data + weight gradient run in passes of <= 4)
    const float* x;                // (N, Cin, H, W)
    const float* w;                // (64, Cin, 3, 3)
    float* y;                      // (N, 64, H, W)
    const float* g_y;
    float* g_x;
    float* partial;                // (blocks, 64 * Cin * 9)
    int ntiles;
    int gpitch, vec16;             // weight gradient: pitch of a g_y row in shared memory; 16-byte copies are aligned
};

// ------------------------------------------------------------------------------------------------ forward
// CIN > 0: the channel loop is unrolled (C_in <= 4, the shapes of the benchmarked step: 104 registers);
// CIN == 0: run-time channel count (5 .. 16 input channels; with the loop unrolled only over the taps)
template <int CIN>
__global__ void __launch_bounds__(kThreads) convin_fwd_kernel(InParams P) {
    extern __shared__ __align__(16) float w_s[];                // [ci * 9 + tap][co]
    const int Cin = CIN > 0 ? CIN : P.Cin;
    for (int i = threadIdx.x; i < 9 * Cin * kCo; i += kThreads) {
        const int k = i / kCo, co = i - k * kCo;
        w_s[i] = __ldg(P.w + co * Cin * 9 + k);
    }
    __syncthreads();
    const int n = blockIdx.y, r0 = blockIdx.x * P.R;
    const int rows = min(P.R, P.H - r0);
    const int t = threadIdx.x;
    if (t >= rows * P.W) return;
    const int r = r0 + t / P.W, c = t % P.W;
    const long long hw = (long long)P.H * P.W;
    float acc[kCo];
#pragma unroll
    for (int co = 0; co < kCo; ++co) acc[co] = 0.f;
    const float* xn = P.x + (long long)n * Cin * hw;
    auto one_channel = [&](int ci) {
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const int rr = r + tap / 3 - 1, cc = c + tap % 3 - 1;
            const float v = (rr >= 0 && rr < P.H && cc >= 0 && cc < P.W) ? __ldg(xn + ci * hw + rr * P.W + cc) : 0.f;
            const float4* wr = reinterpret_cast<const float4*>(w_s + (ci * 9 + tap) * kCo);
#pragma unroll
            for (int q = 0; q < kCo / 4; ++q) {
                const float4 wv = wr[q];
                acc[4 * q] = fmaf(wv.x, v, acc[4 * q]);
                acc[4 * q + 1] = fmaf(wv.y, v, acc[4 * q + 1]);
                acc[4 * q + 2] = fmaf(wv.z, v, acc[4 * q + 2]);
                acc[4 * q + 3] = fmaf(wv.w, v, acc[4 * q + 3]);
            }
        }
    };
    if (CIN > 0) {
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci) one_channel(ci);
    } else {
#pragma unroll 1
        for (int ci = 0; ci < Cin; ++ci) one_channel(ci);
    }
    float* yo = P.y + (long long)n * kCo * hw + r * P.W + c;
#pragma unroll
    for (int co = 0; co < kCo; ++co) yo[co * hw] = acc[co];
}

// ------------------------------------------------------------------------------------------------ data gradient
template <int CIN>
__global__ void __launch_bounds__(kThreads) convin_dgrad_kernel(InParams P) {
    extern __shared__ __align__(16) float dyn[];
    float* w_s = dyn;                                   // [co][CIN * 9] (padded to a multiple of 4)
    constexpr int kK = 9 * CIN, kKp = (kK + 3) & ~3;
    float* z_s = dyn + kCo * kKp;                       // [k][(R + 2) * W]
    for (int i = threadIdx.x; i < kCo * kKp; i += kThreads) {
        const int co = i / kKp, k = i - co * kKp;
        w_s[i] = k < kK ? __ldg(P.w + (co * P.CinT + P.ci0) * 9 + k) : 0.f;
    }
    __syncthreads();
    const int n = blockIdx.y, r0 = blockIdx.x * P.R;
    const int rows = min(P.R, P.H - r0);
    const int zrows = rows + 2, zpix = zrows * P.W;     // rows r0 - 1 .. r0 + rows of z (outside the image: zero)
    const long long hw = (long long)P.H * P.W;
    const int t = threadIdx.x;
    if (t < zpix) {
        const int r = r0 - 1 + t / P.W, c = t % P.W;
        float z[kKp];
#pragma unroll
        for (int k = 0; k < kKp; ++k) z[k] = 0.f;
        if (r >= 0 && r < P.H) {
            const float* g = P.g_y + (long long)n * kCo * hw + r * P.W + c;
#pragma unroll 4
            for (int co = 0; co < kCo; ++co) {
                const float v = __ldg(g + co * hw);
                const float4* wr = reinterpret_cast<const float4*>(w_s + co * kKp);
#pragma unroll
                for (int q = 0; q < kKp / 4; ++q) {
                    const float4 wv = wr[q];
                    z[4 * q] = fmaf(wv.x, v, z[4 * q]);
                    z[4 * q + 1] = fmaf(wv.y, v, z[4 * q + 1]);
                    z[4 * q + 2] = fmaf(wv.z, v, z[4 * q + 2]);
                    z[4 * q + 3] = fmaf(wv.w, v, z[4 * q + 3]);
                }
            }
        }
#pragma unroll
        for (int k = 0; k < kK; ++k) z_s[k * zpix + t] = z[k];
    }
    __syncthreads();
    if (t >= rows * P.W) return;
    const int lr = t / P.W, c = t % P.W;                // local row of the band; its z row is lr + 1
    float* go = P.g_x + ((long long)n * P.CinT + P.ci0) * hw + (r0 + lr) * P.W + c;
#pragma unroll
    for (int ci = 0; ci < CIN; ++ci) {
        float s = 0.f;
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
            const int dy = tap / 3 - 1, dx = tap % 3 - 1;
            const int cc = c - dx;                       // y[p] used x[p + shift]  =>  g_x[q] takes g_y[q - shift]
            if (cc >= 0 && cc < P.W) s += z_s[(ci * 9 + tap) * zpix + (lr + 1 - dy) * P.W + cc];
        }
        go[ci * hw] = s;
    }
}

// ------------------------------------------------------------------------------------------------ weight gradient
__device__ __forceinline__ void cp_async4(float* dst, const float* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src) : "memory");
}

template <int CIN>
__global__ void __launch_bounds__(kThreads) convin_wgrad_kernel(InParams P) {
    extern __shared__ __align__(16) float dyn[];
    const int xw = P.W + 2;
    const int gp = P.gpitch;
    const int gy_elems = kCo * gp, x_elems = CIN * (P.R + 2) * xw;
    // two buffers of { g_y band [co][pixel] (pitch / 4 odd: the 8 channel rows a warp reads hit 8 different banks) |
    // x band [ci][R + 2][W + 2] with a zero border }: the next tile arrives by cp.async while this one is used
    // (16-byte copies; per-element 4-byte copies are processed one lane at a time and were 5x slower)
    float* bufs[2] = {dyn, dyn + gy_elems + x_elems};
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int cog = lane & 7, tg = lane >> 3;            // 8 output channels x tap row dy = tg - 1 (tg == 3: idle)
    float acc[CIN][8][3];
#pragma unroll
    for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int d = 0; d < 3; ++d) acc[ci][j][d] = 0.f;
    const long long hw = (long long)P.H * P.W;
    auto issue = [&](int tile, float* buf) {
        if (tile < P.ntiles) {
            const int n = tile / P.nb, r0 = (tile - n * P.nb) * P.R;
            const int rows = min(P.R, P.H - r0), npix = rows * P.W;
            const float* g = P.g_y + (long long)n * kCo * hw + r0 * P.W;
            if (P.vec16) {
                const int nq = npix >> 2;
                for (int i = threadIdx.x; i < kCo * nq; i += kThreads) {
                    const int co = i / nq, q = i - co * nq;
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(buf + co * gp + q * 4)),
                                 "l"(g + co * hw + q * 4) : "memory");
                }
            } else {
                for (int i = threadIdx.x; i < kCo * npix; i += kThreads) {
                    const int co = i / npix, p = i - co * npix;
                    cp_async4(buf + co * gp + p, g + co * hw + p);
                }
            }
            float* xs = buf + gy_elems;
            const float* xn = P.x + ((long long)n * P.CinT + P.ci0) * hw;
            for (int i = threadIdx.x; i < CIN * (rows + 2) * xw; i += kThreads) {
                const int ci = i / ((rows + 2) * xw), rem = i - ci * (rows + 2) * xw;
                const int rr = r0 - 1 + rem / xw, cc = rem % xw - 1;
                float* dst = xs + ci * (P.R + 2) * xw + rem;
                if (rr >= 0 && rr < P.H && cc >= 0 && cc < P.W) cp_async4(dst, xn + ci * hw + rr * P.W + cc);
                else *dst = 0.f;
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    issue(blockIdx.x, bufs[0]);
    int cur = 0;
    for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x, cur ^= 1) {
        issue(tile + gridDim.x, bufs[cur ^ 1]);
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();
        const int n = tile / P.nb, r0 = (tile - n * P.nb) * P.R;
        const int npix = min(P.R, P.H - r0) * P.W;
        const float* gy_s = bufs[cur];
        const float* x_s = bufs[cur] + gy_elems;
        if (tg < 3) {
            // lane (cog, tg) owns output channels cog + 8 j (conflict-free rows of the transposed tile) and tap row tg
#pragma unroll 2
            for (int p = warp; p < npix; p += kThreads / 32) {
                const int lr = p / P.W, c = p - lr * P.W;
                float gv[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) gv[j] = gy_s[(j * 8 + cog) * gp + p];
#pragma unroll
                for (int ci = 0; ci < CIN; ++ci) {
                    const float* xr = x_s + (ci * (P.R + 2) + lr + tg) * xw + c;      // row lr + dy of the band = tile row lr + 1 + dy
                    const float x0 = xr[0], x1 = xr[1], x2 = xr[2];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        acc[ci][j][0] = fmaf(gv[j], x0, acc[ci][j][0]);
                        acc[ci][j][1] = fmaf(gv[j], x1, acc[ci][j][1]);
                        acc[ci][j][2] = fmaf(gv[j], x2, acc[ci][j][2]);
                    }
                }
            }
        }
        __syncthreads();                                 // buffer cur is free for the tile after next
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
    // block reduction over the 8 warps (fixed order), then this block's partial
    __syncthreads();
    float* red = dyn;                                    // [warp][co][ci][tap]
    if (tg < 3) {
#pragma unroll
        for (int ci = 0; ci < CIN; ++ci)
#pragma unroll
            for (int j = 0; j < 8; ++j)
#pragma unroll
                for (int d = 0; d < 3; ++d) red[warp * kCo * CIN * 9 + ((j * 8 + cog) * CIN + ci) * 9 + tg * 3 + d] = acc[ci][j][d];
    }
    __syncthreads();
    float* part = P.partial + (long long)blockIdx.x * kCo * CIN * 9;
    for (int i = threadIdx.x; i < kCo * CIN * 9; i += kThreads) {
        float s = 0.f;
#pragma unroll
        for (int wv = 0; wv < kThreads / 32; ++wv) s += red[wv * kCo * CIN * 9 + i];
        part[i] = s;
    }
}

// partial[b][co][ci of the pass][tap] summed over the blocks in order -> g_w[co][ci0 + ci][tap] of the full tensor
__global__ void convin_reduce_kernel(int nparts, int cin, int cint, int ci0, const float* __restrict__ partial, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, n = kCo * cin * 9;
    if (i >= n) return;
    float s = 0.f;
    for (int b = 0; b < nparts; ++b) s += partial[(long long)b * n + i];
    const int co = i / (cin * 9), rem = i - co * cin * 9;
    out[(co * cint + ci0) * 9 + rem] = s;
}

int fill(InParams& P, const char* who, int N, int Cin, int H, int W) {
    CMF_REQUIRE(N >= 0, "%s: N = %d", who, N);
    CMF_REQUIRE(Cin >= 1 && Cin <= kMaxCin && H >= 1 && W >= 1 && W <= 64 && H <= 256, "%s: C_in = %d (1..%d), H = %d, W = %d (W <= 64)", who,
                Cin, kMaxCin, H, W);
    P.N = N; P.Cin = Cin; P.CinT = Cin; P.ci0 = 0; P.H = H; P.W = W;
    return 0;
}

int bands(int H, int W, int max_pix, int halo, int* R) {
    int r = max_pix / W - halo;                      // most rows a block can hold
    if (r > H) r = H;
    if (r < 1) r = 1;
    const int nb = (int)ceil_div(H, r);
    *R = (int)ceil_div(H, nb);                       // balanced bands
    return nb;
}

template <int CIN> size_t dgrad_smem(int R, int W) { return (size_t)(kCo * ((9 * CIN + 3) & ~3) + 9 * CIN * (R + 2) * W) * sizeof(float); }
int wgrad_pitch(int R, int W) {
    int p = (R * W + 3) & ~3;
    if (((p >> 2) & 1) == 0) p += 4;                 // pitch / 4 odd
    return p;
}
// bands of the weight gradient: R W a multiple of 4 where the width allows it (16-byte copies)
int wgrad_bands(int H, int W, int* R) {
    int nb = bands(H, W, 128, 0, R);
    if ((*R * W) % 4 != 0 && W % 2 == 0) {
        const int r = (*R + 1) * W <= 128 ? *R + 1 : *R - 1;
        if (r >= 1 && (r * W) % 4 == 0) { *R = r; nb = (int)ceil_div(H, r); }
    }
    return nb;
}
size_t wgrad_smem(int Cin, int R, int W) {
    const size_t tiles = 2 * ((size_t)kCo * wgrad_pitch(R, W) + (size_t)Cin * (R + 2) * (W + 2));
    const size_t red = (size_t)(kThreads / 32) * kCo * Cin * 9;
    return (tiles > red ? tiles : red) * sizeof(float);
}

DeviceOnce g_convin_attr;
int set_attrs() {
    if (g_convin_attr.done()) return 0;
    cudaError_t e = cudaSuccess;
    const int big = 96 * 1024;
#define CMF_ATTR(k) if (e == cudaSuccess) e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, big)
    CMF_ATTR(convin_dgrad_kernel<1>); CMF_ATTR(convin_dgrad_kernel<2>); CMF_ATTR(convin_dgrad_kernel<3>); CMF_ATTR(convin_dgrad_kernel<4>);
    CMF_ATTR(convin_wgrad_kernel<1>); CMF_ATTR(convin_wgrad_kernel<2>); CMF_ATTR(convin_wgrad_kernel<3>); CMF_ATTR(convin_wgrad_kernel<4>);
#undef CMF_ATTR
    CMF_REQUIRE(e == cudaSuccess, "convin: %s", cudaGetErrorString(e));
    g_convin_attr.mark();
    return 0;
}

}  // namespace
}  // namespace cmf

using namespace cmf;

#define CMF_CIN_DISPATCH(CALL)          \
    switch (P.Cin) {                    \
        case 1: CALL(1); break;         \
        case 2: CALL(2); break;         \
        case 3: CALL(3); break;         \
        default: CALL(4); break;        \
    }

extern "C" {

int cmf_convin_usable(int Cin, int Cout, int H, int W) {
    return Cout == kCo && Cin >= 1 && Cin <= kMaxCin && H >= 1 && H <= 256 && W >= 1 && W <= 64;
}

int cmf_convin_wgrad_partials(int N, int H, int W) {
    int R;
    const long long tiles = (long long)N * wgrad_bands(H, W, &R);
    return (int)(tiles < 3 * kNumSMs ? tiles : 3 * kNumSMs);      // three 60 KB blocks are resident per SM
}

int cmf_convin_fwd(int N, int Cin, int H, int W, const float* x, const float* w, float* y, cmf_stream_t stream) {
    InParams P{};
    if (int rc = fill(P, "convin_fwd", N, Cin, H, W)) return rc;
    if (N == 0) return 0;
    CMF_REQUIRE(x && w && y, "convin_fwd: null pointer");
    CMF_REQUIRE(N <= 65535, "convin_fwd: N = %d exceeds the grid (split the batch)", N);
    P.x = x; P.w = w; P.y = y;
    P.nb = bands(H, W, kThreads, 0, &P.R);
    const dim3 grid(P.nb, N);
    const size_t smem = (size_t)9 * Cin * kCo * sizeof(float);
    cudaStream_t st = (cudaStream_t)stream;
    switch (Cin) {
        case 1: convin_fwd_kernel<1><<<grid, kThreads, smem, st>>>(P); break;
        case 2: convin_fwd_kernel<2><<<grid, kThreads, smem, st>>>(P); break;
        case 3: convin_fwd_kernel<3><<<grid, kThreads, smem, st>>>(P); break;
        case 4: convin_fwd_kernel<4><<<grid, kThreads, smem, st>>>(P); break;
        default: convin_fwd_kernel<0><<<grid, kThreads, smem, st>>>(P); break;
    }
    CMF_CHECK_LAUNCH();
    return 0;
}

int cmf_convin_dgrad(int N, int Cin, int H, int W, const float* g_y, const float* w, float* g_x, cmf_stream_t stream) {
    InParams P{};
    if (int rc = fill(P, "convin_dgrad", N, Cin, H, W)) return rc;
    if (N == 0) return 0;
    CMF_REQUIRE(g_y && w && g_x, "convin_dgrad: null pointer");
    CMF_REQUIRE(N <= 65535, "convin_dgrad: N = %d exceeds the grid (split the batch)", N);
    if (int rc = set_attrs()) return rc;
    P.g_y = g_y; P.w = w; P.g_x = g_x;
    P.nb = bands(H, W, kThreads, 2, &P.R);
    const dim3 grid(P.nb, N);
    for (int ci0 = 0; ci0 < Cin; ci0 += kChunk) {           // input channels are independent: passes of <= 4
        P.CinT = Cin; P.ci0 = ci0; P.Cin = Cin - ci0 < kChunk ? Cin - ci0 : kChunk;
#define CALL(C) convin_dgrad_kernel<C><<<grid, kThreads, dgrad_smem<C>(P.R, W), (cudaStream_t)stream>>>(P)
        CMF_CIN_DISPATCH(CALL)
#undef CALL
        CMF_CHECK_LAUNCH();
    }
    return 0;
}

int cmf_convin_wgrad(int N, int Cin, int H, int W, const float* x, const float* g_y, float* partial, float* g_w, cmf_stream_t stream) {
    InParams P{};
    if (int rc = fill(P, "convin_wgrad", N, Cin, H, W)) return rc;
    CMF_REQUIRE(g_w, "convin_wgrad: null gradient");
    const int nparts = cmf_convin_wgrad_partials(N, H, W);
    if (N > 0) {
        CMF_REQUIRE(x && g_y && partial, "convin_wgrad: null pointer");
        if (int rc = set_attrs()) return rc;
        P.x = x; P.g_y = g_y; P.partial = partial;
        P.nb = wgrad_bands(H, W, &P.R);
        P.ntiles = N * P.nb;
        P.gpitch = wgrad_pitch(P.R, W);
        const int last = H - (P.nb - 1) * P.R;           // rows of the last band
        P.vec16 = ((long long)H * W) % 4 == 0 && (P.R * W) % 4 == 0 && (last * W) % 4 == 0 && aligned16(g_y);
    }
    for (int ci0 = 0; ci0 < Cin; ci0 += kChunk) {           // input channels are independent: passes of <= 4
        P.CinT = Cin; P.ci0 = ci0; P.Cin = Cin - ci0 < kChunk ? Cin - ci0 : kChunk;
        if (N > 0) {
            const size_t smem = wgrad_smem(P.Cin, P.R, W);
            CMF_REQUIRE(smem <= 96 * 1024, "convin_wgrad: %zu bytes of shared memory", smem);
#define CALL(C) convin_wgrad_kernel<C><<<nparts, kThreads, smem, (cudaStream_t)stream>>>(P)
            CMF_CIN_DISPATCH(CALL)
#undef CALL
            CMF_CHECK_LAUNCH();
        }
        const int n = kCo * P.Cin * 9;
        convin_reduce_kernel<<<(unsigned)ceil_div(n, 128), 128, 0, (cudaStream_t)stream>>>(nparts, P.Cin, Cin, ci0, partial, g_w);
        CMF_CHECK_LAUNCH();
    }
    return 0;
}

}  // extern "C"
