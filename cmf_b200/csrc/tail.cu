// Tail of the image couplers' conditioner (ReLU -> 1x1 convolution -> scaled tanh), primal + d tangent
// images, evaluated only at the spatial positions the coupling layer uses, in one launch; one launch
// (+ a fixed-order reduction) for its adjoint.
//
// Replaces (k-flouris/cmf @ df4521aa):
//   cmf/models/components/networks.py:156-161   the ReLU and the 1x1 output convolution of get_resnet
//   cmf/models/components/networks.py:96-113    ScaledTanh2dModule (weights * tanh(.) + bias) and its jvp
//   cmf/models/components/jvp_layers.py:38-64   relu-JVP mask, bias-free convolution on the tangent
// and the gather of the used positions (`acl.py:53-66` multiplies everything else by zero).  In the
// library path this was a gather kernel, two broadcast mask products over the (B d, C, P) tangents, a
// cuBLAS SIMT GEMM with an O x C weight (O = 2..24 output channels), tanh and two more products.
//
// HBM-bound: every hidden value of the used positions is read once, (1 + d) C P floats per sample.
// A thread owns one spatial position of one sample: it rebuilds the primal's ReLU pattern (C <= 64 sign
// bits), its tanh and tanh', then walks the d tangent images.  Loads / stores are coalesced over the
// positions of a row.  The adjoint writes DENSE gradients of the hidden tensors (zeros at the unused
// positions, so no memset pass) and reduces the O x C weight gradient through shared memory per block,
// then over the blocks in block order (bitwise reproducible).
#include "common.cuh"

namespace cmf {
namespace {

constexpr int kTailThreads = 128;
constexpr int kTailC = CMF_TAIL_MAX_C;          // 64
constexpr int kVPitch = kTailThreads + 1;       // transposed staging tile of the weight-gradient reduction
constexpr int kTailChunk = 8;                   // tangent images per block: (HW / 128) x B x (d / 8) blocks fill the GPU at B = 50

struct TailParams {
    int B, d, C, HW, P, O;
    const float* h;        // (B, C, HW) primal hidden activations before the ReLU
    const float* Hv;       // (B d, C, HW) tangents of the same
    const int* inv;        // (HW) position -> compact index, < 0: unused; null: identity (P == HW)
    const float* w2;       // (O, C)
    const float* b2;       // (O)
    const float* sw;       // (O) scale of the tanh
    const float* sb;       // (O) shift of the tanh
    float* y;              // (B, O, P)
    float* Y;              // (B, d, O, P)
    const float* g_y;      // adjoint inputs (g_y may be null)
    const float* g_Y;
    float* g_h;            // (B, C, HW) dense
    float* g_Hv;           // (B d, C, HW) dense
    float* partial;        // (blocks of both adjoint launches, O * C + 3 * O)
    float* gdy;            // (chunks, B, O, P): per tangent chunk, sum_k g_Y_k (W2 (Hv_k * mask)) -- scratch of the adjoint
    int kc, nchunks;       // tangents per block (blockIdx.z = chunk)
};

// primal at one position: ReLU pattern, th = tanh(W relu(h) + b)
template <int OMAX>
__device__ __forceinline__ void primal_at(const TailParams& P, const float* w_s, const float* hb, int q, unsigned long long& bits,
                                          float (&th)[OMAX]) {
    float pre[OMAX];
#pragma unroll
    for (int o = 0; o < OMAX; ++o) pre[o] = 0.f;
    bits = 0ull;
#pragma unroll 8
    for (int c = 0; c < P.C; ++c) {
        const float v = __ldg(hb + (long long)c * P.HW + q);
        if (v > 0.f) {
            bits |= 1ull << c;
#pragma unroll
            for (int o = 0; o < OMAX; ++o) pre[o] = fmaf(w_s[c * OMAX + o], v, pre[o]);
        }
    }
#pragma unroll
    for (int o = 0; o < OMAX; ++o) th[o] = o < P.O ? tanhf(pre[o] + __ldg(P.b2 + o)) : 0.f;
}

__device__ __forceinline__ void stage_weights(const TailParams& P, float* w_s, int omax) {
    for (int i = threadIdx.x; i < kTailC * omax; i += kTailThreads) {
        const int c = i / omax, o = i - c * omax;
        w_s[i] = (c < P.C && o < P.O) ? __ldg(P.w2 + o * P.C + c) : 0.f;
    }
}

template <int OMAX>
__global__ void __launch_bounds__(kTailThreads) tail_fwd_kernel(TailParams P) {
    __shared__ float w_s[kTailC * OMAX];       // [c][o]
    stage_weights(P, w_s, OMAX);
    __syncthreads();
    const int q = blockIdx.x * kTailThreads + threadIdx.x, b = blockIdx.y;
    if (q >= P.HW) return;
    const int p = P.inv ? __ldg(P.inv + q) : q;
    if (p < 0) return;
    unsigned long long bits;
    float th[OMAX];
    primal_at<OMAX>(P, w_s, P.h + (long long)b * P.C * P.HW, q, bits, th);
    float dy[OMAX];
#pragma unroll
    for (int o = 0; o < OMAX; ++o) {
        const float sw = o < P.O ? __ldg(P.sw + o) : 0.f;
        dy[o] = (1.f - th[o] * th[o]) * sw;
        if (o < P.O && blockIdx.z == 0) P.y[((long long)b * P.O + o) * P.P + p] = sw * th[o] + __ldg(P.sb + o);
    }
    const int k0 = blockIdx.z * P.kc, k1 = min(P.d, k0 + P.kc);
    for (int k = k0; k < k1; ++k) {
        const float* src = P.Hv + (long long)(b * P.d + k) * P.C * P.HW + q;
        float acc[OMAX];
#pragma unroll
        for (int o = 0; o < OMAX; ++o) acc[o] = 0.f;
        if (P.C == kTailC) {
            // all 64 loads of the tangent image in flight before the first use (the kernel is bound by memory
            // latency: ncu showed 1.15 TB/s with 16 loads in flight per thread)
            float v[kTailC];
#pragma unroll
            for (int c = 0; c < kTailC; ++c) v[c] = ((bits >> c) & 1ull) ? __ldcs(src + (long long)c * P.HW) : 0.f;
#pragma unroll
            for (int c = 0; c < kTailC; ++c)
#pragma unroll
                for (int o = 0; o < OMAX; ++o) acc[o] = fmaf(w_s[c * OMAX + o], v[c], acc[o]);
        } else {
#pragma unroll 8
            for (int c = 0; c < P.C; ++c) {
                const float v = ((bits >> c) & 1ull) ? __ldcs(src + (long long)c * P.HW) : 0.f;
#pragma unroll
                for (int o = 0; o < OMAX; ++o) acc[o] = fmaf(w_s[c * OMAX + o], v, acc[o]);
            }
        }
        float* dst = P.Y + ((long long)(b * P.d + k) * P.O) * P.P + p;
#pragma unroll
        for (int o = 0; o < OMAX; ++o)
            if (o < P.O) dst[(long long)o * P.P] = acc[o] * dy[o];
    }
}

// block-wide: part[o * C + c] += sum over the block's threads of z[o] * v[c], through the staging tile
template <int OMAX>
__device__ __forceinline__ void reduce_outer(const TailParams& P, float* v_s, float* z_s, float (&wacc)[(OMAX * kTailC + kTailThreads - 1) / kTailThreads]) {
    __syncthreads();                                           // v_s / z_s of this round are complete
    constexpr int kPairs = (OMAX * kTailC + kTailThreads - 1) / kTailThreads;
#pragma unroll
    for (int i = 0; i < kPairs; ++i) {
        const int pair = threadIdx.x + i * kTailThreads;       // pair = o * 64 + c
        const int o = pair / kTailC, c = pair - o * kTailC;
        if (o < P.O && c < P.C) {
            const float* vr = v_s + c * kVPitch;
            const float* zr = z_s + o * kTailThreads;
            float s = 0.f;
#pragma unroll 8
            for (int t = 0; t < kTailThreads; ++t) s = fmaf(zr[t], vr[t], s);
            wacc[i] += s;
        }
    }
    __syncthreads();                                           // the tile may be overwritten
}

// adjoint, tangent part: one block = 128 positions x one sample x one chunk of tangents
template <int OMAX>
__global__ void __launch_bounds__(kTailThreads) tail_bwd_tangent_kernel(TailParams P) {
    extern __shared__ float dyn[];
    float* w_s = dyn;                                   // [c][o]
    float* v_s = w_s + kTailC * OMAX;                   // [c][thread], pitch kVPitch
    float* z_s = v_s + kTailC * kVPitch;                // [o][thread]
    stage_weights(P, w_s, OMAX);
    __syncthreads();
    constexpr int kPairs = (OMAX * kTailC + kTailThreads - 1) / kTailThreads;
    float wacc[kPairs];
#pragma unroll
    for (int i = 0; i < kPairs; ++i) wacc[i] = 0.f;
    const int t = threadIdx.x;
    const int q = blockIdx.x * kTailThreads + t, b = blockIdx.y;
    const int p = q < P.HW ? (P.inv ? __ldg(P.inv + q) : q) : -1;
    const bool on = p >= 0;                              // this thread's position is used by the coupling layer
    const bool inside = q < P.HW;
    unsigned long long bits = 0ull;
    float th[OMAX], dy[OMAX], gdy[OMAX];
#pragma unroll
    for (int o = 0; o < OMAX; ++o) { th[o] = 0.f; gdy[o] = 0.f; }
    if (on) primal_at<OMAX>(P, w_s, P.h + (long long)b * P.C * P.HW, q, bits, th);
#pragma unroll
    for (int o = 0; o < OMAX; ++o) dy[o] = o < P.O ? (1.f - th[o] * th[o]) * __ldg(P.sw + o) : 0.f;
    const int k0 = blockIdx.z * P.kc, k1 = min(P.d, k0 + P.kc);
    for (int k = k0; k < k1; ++k) {
        const long long n = (long long)b * P.d + k;
        float gz[OMAX], gY[OMAX], acc[OMAX];
#pragma unroll
        for (int o = 0; o < OMAX; ++o) {
            gY[o] = (on && o < P.O) ? __ldcs(P.g_Y + (n * P.O + o) * P.P + p) : 0.f;
            gz[o] = gY[o] * dy[o];
            acc[o] = 0.f;
            z_s[o * kTailThreads + t] = gz[o];
        }
        const float* src = P.Hv + n * P.C * P.HW + q;
        float* dst = P.g_Hv + n * P.C * P.HW + q;
        if (P.C == kTailC) {
            // every load of the tangent image in flight before the first use (memory-latency bound otherwise)
            float v[kTailC];
#pragma unroll
            for (int c = 0; c < kTailC; ++c) v[c] = (on && ((bits >> c) & 1ull)) ? __ldcs(src + (long long)c * P.HW) : 0.f;
#pragma unroll
            for (int c = 0; c < kTailC; ++c) {
                const bool live = on && ((bits >> c) & 1ull);
                float g = 0.f;
#pragma unroll
                for (int o = 0; o < OMAX; ++o) {
                    acc[o] = fmaf(w_s[c * OMAX + o], v[c], acc[o]);
                    g = fmaf(w_s[c * OMAX + o], gz[o], g);
                }
                v_s[c * kVPitch + t] = v[c];
                if (inside) __stcs(dst + (long long)c * P.HW, live ? g : 0.f);
            }
        } else {
#pragma unroll 8
            for (int c = 0; c < P.C; ++c) {
                const bool live = on && ((bits >> c) & 1ull);
                const float v = live ? __ldcs(src + (long long)c * P.HW) : 0.f;
                float g = 0.f;
#pragma unroll
                for (int o = 0; o < OMAX; ++o) {
                    acc[o] = fmaf(w_s[c * OMAX + o], v, acc[o]);
                    g = fmaf(w_s[c * OMAX + o], gz[o], g);
                }
                v_s[c * kVPitch + t] = v;
                if (inside) __stcs(dst + (long long)c * P.HW, live ? g : 0.f);
            }
        }
#pragma unroll
        for (int o = 0; o < OMAX; ++o) gdy[o] = fmaf(gY[o], acc[o], gdy[o]);
        reduce_outer<OMAX>(P, v_s, z_s, wacc);
    }
    if (on) {
#pragma unroll
        for (int o = 0; o < OMAX; ++o)
            if (o < P.O) P.gdy[(((long long)blockIdx.z * P.B + b) * P.O + o) * P.P + p] = gdy[o];
    }
    const int n = P.O * P.C + 3 * P.O;
    float* part = P.partial + (((long long)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * n;
#pragma unroll
    for (int i = 0; i < kPairs; ++i) {
        const int pair = t + i * kTailThreads;
        const int o = pair / kTailC, c = pair - o * kTailC;
        if (o < P.O && c < P.C) part[o * P.C + c] = wacc[i];
    }
    if (t < 3 * P.O) part[P.O * P.C + t] = 0.f;
}

// adjoint, primal part (after the tangent launch): one block = 128 positions x one sample
template <int OMAX>
__global__ void __launch_bounds__(kTailThreads) tail_bwd_primal_kernel(TailParams P, int first_block) {
    extern __shared__ float dyn[];
    float* w_s = dyn;
    float* v_s = w_s + kTailC * OMAX;
    float* z_s = v_s + kTailC * kVPitch;
    float* r_s = z_s + OMAX * kTailThreads;             // [3 * OMAX][warps] block sums of the small gradients
    stage_weights(P, w_s, OMAX);
    __syncthreads();
    constexpr int kPairs = (OMAX * kTailC + kTailThreads - 1) / kTailThreads;
    float wacc[kPairs];
#pragma unroll
    for (int i = 0; i < kPairs; ++i) wacc[i] = 0.f;
    const int t = threadIdx.x;
    const int q = blockIdx.x * kTailThreads + t, b = blockIdx.y;
    const int p = q < P.HW ? (P.inv ? __ldg(P.inv + q) : q) : -1;
    const bool on = p >= 0, inside = q < P.HW;
    unsigned long long bits = 0ull;
    float th[OMAX];
#pragma unroll
    for (int o = 0; o < OMAX; ++o) th[o] = 0.f;
    if (on) primal_at<OMAX>(P, w_s, P.h + (long long)b * P.C * P.HW, q, bits, th);
    // y = sw th + sb, dy = (1 - th^2) sw;  gdy = sum over the tangent chunks, in chunk order
    float gpre[OMAX], gsw[OMAX], gsb[OMAX];
#pragma unroll
    for (int o = 0; o < OMAX; ++o) {
        float gdy = 0.f;
        if (on && o < P.O && P.d > 0)
            for (int ch = 0; ch < P.nchunks; ++ch) gdy += __ldg(P.gdy + (((long long)ch * P.B + b) * P.O + o) * P.P + p);
        const float sw = o < P.O ? __ldg(P.sw + o) : 0.f;
        const float gy = (on && o < P.O && P.g_y) ? __ldg(P.g_y + ((long long)b * P.O + o) * P.P + p) : 0.f;
        const float one = 1.f - th[o] * th[o];
        gsb[o] = gy;
        gsw[o] = gy * th[o] + gdy * one;
        gpre[o] = (gy * sw - 2.f * th[o] * sw * gdy) * one;
        z_s[o * kTailThreads + t] = gpre[o];
    }
    {
        const float* src = P.h + (long long)b * P.C * P.HW + q;
        float* dst = P.g_h + (long long)b * P.C * P.HW + q;
#pragma unroll 8
        for (int c = 0; c < P.C; ++c) {
            const bool live = on && ((bits >> c) & 1ull);
            float g = 0.f;
#pragma unroll
            for (int o = 0; o < OMAX; ++o) g = fmaf(w_s[c * OMAX + o], gpre[o], g);
            v_s[c * kVPitch + t] = live ? __ldg(src + (long long)c * P.HW) : 0.f;
            if (inside) dst[(long long)c * P.HW] = live ? g : 0.f;
        }
    }
    reduce_outer<OMAX>(P, v_s, z_s, wacc);
    // block sums of g_b2 (= sum gpre), g_sw, g_sb: warp sums, then the warps in order
    const int warp = t >> 5, lane = t & 31;
#pragma unroll
    for (int o = 0; o < OMAX; ++o) {
        const float a = warp_sum(gpre[o]), bb = warp_sum(gsw[o]), cc = warp_sum(gsb[o]);
        if (lane == 0) {
            r_s[(0 * OMAX + o) * 4 + warp] = a;
            r_s[(1 * OMAX + o) * 4 + warp] = bb;
            r_s[(2 * OMAX + o) * 4 + warp] = cc;
        }
    }
    __syncthreads();
    const int n = P.O * P.C + 3 * P.O;
    float* part = P.partial + ((long long)first_block + (long long)blockIdx.y * gridDim.x + blockIdx.x) * n;
#pragma unroll
    for (int i = 0; i < kPairs; ++i) {
        const int pair = t + i * kTailThreads;
        const int o = pair / kTailC, c = pair - o * kTailC;
        if (o < P.O && c < P.C) part[o * P.C + c] = wacc[i];
    }
    if (t < 3 * P.O) {
        const int which = t / P.O, o = t - which * P.O;
        const float* r = r_s + (which * OMAX + o) * 4;
        part[P.O * P.C + t] = ((r[0] + r[1]) + r[2]) + r[3];
    }
}

// out[i] = sum over the blocks, in block order
// one block per output element: 128 threads take the partials b = t, t + 128, ... in order, then a fixed tree
__global__ void tail_reduce_kernel(int nblocks, int n, const float* __restrict__ partial, float* __restrict__ out) {
    __shared__ float r[128];
    const int i = blockIdx.x, t = threadIdx.x;
    float s = 0.f;
    for (int b = t; b < nblocks; b += 128) s += partial[(long long)b * n + i];
    r[t] = s;
    __syncthreads();
    for (int w = 64; w > 0; w >>= 1) {
        if (t < w) r[t] += r[t + w];
        __syncthreads();
    }
    if (t == 0) out[i] = r[0];
}

int check(const char* who, int B, int d, int C, int HW, int P, int O) {
    CMF_REQUIRE(B >= 0 && d >= 0 && HW >= 1 && P >= 0 && P <= HW, "%s: B = %d, d = %d, HW = %d, P = %d", who, B, d, HW, P);
    CMF_REQUIRE(C >= 1 && C <= kTailC && O >= 1 && O <= CMF_TAIL_MAX_OUT, "%s: C = %d (max %d), O = %d (max %d)", who, C, kTailC, O, CMF_TAIL_MAX_OUT);
    CMF_REQUIRE((long long)B * (d > 0 ? d : 1) * C * HW < (1ll << 40), "%s: tensor too large", who);
    return 0;
}

template <int OMAX>
size_t bwd_smem() { return (size_t)(kTailC * OMAX + kTailC * kVPitch + OMAX * kTailThreads + 3 * OMAX * 4) * sizeof(float); }

int chunks_of(int d) { return d > 0 ? (int)ceil_div(d, kTailChunk) : 1; }

DeviceOnce g_tail_attr;

}  // namespace
}  // namespace cmf

using namespace cmf;

extern "C" {

int cmf_resnet_tail_usable(int C, int O) { return C >= 1 && C <= kTailC && O >= 1 && O <= CMF_TAIL_MAX_OUT; }

long long cmf_resnet_tail_partial_floats(int B, int d, int HW, int P, int C, int O) {
    const long long blocks = (long long)B * ceil_div(HW, kTailThreads);
    const long long nch = d > 0 ? chunks_of(d) : 0;
    return blocks * (nch + 1) * ((long long)O * C + 3 * O) + nch * B * O * P;
}

int cmf_resnet_tail_fwd(int B, int d, int C, int HW, int P, int O, const float* h, const float* Hv, const int32_t* inv,
                        const float* w2, const float* b2, const float* sw, const float* sb, float* y, float* Y,
                        cmf_stream_t stream) {
    if (int rc = check("resnet_tail_fwd", B, d, C, HW, P, O)) return rc;
    if (B == 0 || P == 0) return 0;
    CMF_REQUIRE(h && w2 && b2 && sw && sb && y && (d == 0 || (Hv && Y)), "resnet_tail_fwd: null pointer");
    CMF_REQUIRE(inv || P == HW, "resnet_tail_fwd: P != HW needs the position map");
    CMF_REQUIRE(B <= 65535, "resnet_tail_fwd: B = %d exceeds the grid", B);
    TailParams T{};
    T.B = B; T.d = d; T.C = C; T.HW = HW; T.P = P; T.O = O;
    T.h = h; T.Hv = Hv; T.inv = inv; T.w2 = w2; T.b2 = b2; T.sw = sw; T.sb = sb; T.y = y; T.Y = Y;
    T.kc = kTailChunk; T.nchunks = chunks_of(d);
    CMF_REQUIRE(T.nchunks <= 65535, "resnet_tail_fwd: d = %d exceeds the grid", d);
    const dim3 grid((unsigned)ceil_div(HW, kTailThreads), B, T.nchunks);
    if (O <= 4) tail_fwd_kernel<4><<<grid, kTailThreads, 0, (cudaStream_t)stream>>>(T);
    else if (O <= 8) tail_fwd_kernel<8><<<grid, kTailThreads, 0, (cudaStream_t)stream>>>(T);
    else if (O <= 16) tail_fwd_kernel<16><<<grid, kTailThreads, 0, (cudaStream_t)stream>>>(T);
    else tail_fwd_kernel<32><<<grid, kTailThreads, 0, (cudaStream_t)stream>>>(T);
    CMF_CHECK_LAUNCH();
    return 0;
}

int cmf_resnet_tail_bwd(int B, int d, int C, int HW, int P, int O, const float* h, const float* Hv, const int32_t* inv,
                        const float* w2, const float* b2, const float* sw, const float* sb, const float* g_y,
                        const float* g_Y, float* g_h, float* g_Hv, float* partial, float* g_params, cmf_stream_t stream) {
    if (int rc = check("resnet_tail_bwd", B, d, C, HW, P, O)) return rc;
    CMF_REQUIRE(g_params, "resnet_tail_bwd: null gradient buffer");
    const int blocks = (int)(B * ceil_div(HW, kTailThreads));
    const int nch = d > 0 ? chunks_of(d) : 0;
    const int nblocks = blocks * (nch + 1);
    const int n = O * C + 3 * O;
    if (B > 0) {
        CMF_REQUIRE(h && w2 && b2 && sw && sb && g_h && partial && (d == 0 || (Hv && g_Y && g_Hv)), "resnet_tail_bwd: null pointer");
        CMF_REQUIRE(inv || P == HW, "resnet_tail_bwd: P != HW needs the position map");
        CMF_REQUIRE(B <= 65535 && nch <= 65535, "resnet_tail_bwd: B = %d, d = %d exceed the grid", B, d);
        TailParams T{};
        T.B = B; T.d = d; T.C = C; T.HW = HW; T.P = P; T.O = O;
        T.h = h; T.Hv = Hv; T.inv = inv; T.w2 = w2; T.b2 = b2; T.sw = sw; T.sb = sb;
        T.g_y = g_y; T.g_Y = g_Y; T.g_h = g_h; T.g_Hv = g_Hv; T.partial = partial;
        T.gdy = partial + (long long)nblocks * n;
        T.kc = kTailChunk; T.nchunks = nch;
        if (!g_tail_attr.done()) {
            cudaError_t e = cudaFuncSetAttribute(tail_bwd_tangent_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd_smem<32>());
            if (e == cudaSuccess) e = cudaFuncSetAttribute(tail_bwd_primal_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bwd_smem<32>());
            CMF_REQUIRE(e == cudaSuccess, "resnet_tail_bwd: %s", cudaGetErrorString(e));
            g_tail_attr.mark();
        }
        const dim3 grid((unsigned)ceil_div(HW, kTailThreads), B, nch > 0 ? nch : 1);
        const dim3 gridp((unsigned)ceil_div(HW, kTailThreads), B);
        const int first = blocks * nch;
        cudaStream_t st = (cudaStream_t)stream;
#define CMF_TAIL_BWD(OM)                                                                              \
        do {                                                                                          \
            if (nch > 0) {                                                                            \
                tail_bwd_tangent_kernel<OM><<<grid, kTailThreads, bwd_smem<OM>(), st>>>(T);           \
                CMF_CHECK_LAUNCH();                                                                   \
            }                                                                                         \
            tail_bwd_primal_kernel<OM><<<gridp, kTailThreads, bwd_smem<OM>(), st>>>(T, first);        \
            CMF_CHECK_LAUNCH();                                                                       \
        } while (0)
        if (O <= 4) CMF_TAIL_BWD(4);
        else if (O <= 8) CMF_TAIL_BWD(8);
        else if (O <= 16) CMF_TAIL_BWD(16);
        else CMF_TAIL_BWD(32);
#undef CMF_TAIL_BWD
    }
    tail_reduce_kernel<<<n, 128, 0, (cudaStream_t)stream>>>(nblocks, n, partial, g_params);
    CMF_CHECK_LAUNCH();
    return 0;
}

}  // extern "C"
