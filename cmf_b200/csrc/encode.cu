// Encode-pass and loss-epilogue kernels: the elementwise / per-sample-reduction work
// around the head that the reference issues as strings of small ATen ops.
//
//   preproc_logit        x -> logit(mult * (x + noise) + add) with its log-Jacobian per sample:
//                        DequantizationDensity (wrapper.py:27-30) + scalar-mult + scalar-add + logit
//                        (math.py:41-53,70-105; schemas.py:30-50), four modules, ~12 launches -> 1
//   coupling_x_to_z      affine coupling in the ENCODE direction, z_m = (x_m + t) e^{s}, log-jac = sum s,
//                        with the passthrough half copied through (acl.py:43-46,101-111): split, exp,
//                        add, mul, combine / index_put, masked sum -> 1 (and 1 for the adjoint)
//   sumsq_diff           out[b] = sum_i (a[b,i] - c[b,i])^2 (c may be null): the reconstruction error
//                        (non_square.py:111-114) and the Gaussian log-prob's quadratic form
//                        (gaussian.py:9-22), + adjoint
//
// All HBM-bound, primal-sized (B x D, no tangent factor): one CTA per sample, 128-bit accesses are not
// worth their alignment constraints at these sizes (D <= 3072).  Algorithmic bytes: 12 D per sample
// (preproc: read x, noise, write y), 16 D (coupling forward), 8 D (sumsq).
#include "common.cuh"

namespace cmf {
namespace {

constexpr int kEncThreads = 256;

__device__ __forceinline__ float block_sum(float v, float* red) {
    v = warp_sum(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = (threadIdx.x < kEncThreads / 32) ? red[threadIdx.x] : 0.f;
    if (warp == 0) {
        t = warp_sum(t);
        if (lane == 0) red[0] = t;
    }
    __syncthreads();
    const float out = red[0];
    __syncthreads();
    return out;
}

__global__ void __launch_bounds__(kEncThreads)
preproc_logit_kernel(int n, float* __restrict__ x, const float* __restrict__ noise, float mult, float add, float lj0,
                     float* __restrict__ y, float* __restrict__ logjac) {
    __shared__ float red[kEncThreads / 32];
    const size_t base = (size_t)blockIdx.x * n;
    float acc = 0.f;
    for (int i = threadIdx.x; i < n; i += kEncThreads) {
        float v = x[base + i];
        if (noise) {                       // the reference adds the noise IN PLACE to the caller's batch
            v += noise[base + i];
            x[base + i] = v;
        }
        const float u = mult * v + add;
        y[base + i] = logf(u) - logf(1.f - u);                       // math.py:47
        const float uc = fminf(fmaxf(u, 1e-7f), 1.f - 1e-7f);       // math.py:50-52 (clamped in the log-Jacobian only)
        acc += -logf(uc) - logf(1.f - uc);
    }
    const float tot = block_sum(acc, red);
    if (threadIdx.x == 0) logjac[blockIdx.x] = lj0 + tot;
}

// pos2j (D): compact index j of the modified element at flat position i, or -1 (passthrough)
// qmap (M) or null: conditioner position q of compact element j (null: q = j)
__global__ void __launch_bounds__(kEncThreads)
coupling_x_to_z_fwd_kernel(int D, int Q, const float* __restrict__ x, const float* __restrict__ st,
                           const int32_t* __restrict__ pos2j, const int32_t* __restrict__ qmap,
                           float* __restrict__ z, float* __restrict__ logjac) {
    __shared__ float red[kEncThreads / 32];
    const size_t xb = (size_t)blockIdx.x * D, sb = (size_t)blockIdx.x * 2 * Q;
    float acc = 0.f;
    for (int i = threadIdx.x; i < D; i += kEncThreads) {
        const int j = __ldg(pos2j + i);
        float v = x[xb + i];
        if (j >= 0) {
            const int q = qmap ? __ldg(qmap + j) : j;
            const float s = st[sb + Q + q];
            v = (v + st[sb + q]) * expf(s);
            acc += s;
        }
        z[xb + i] = v;
    }
    const float tot = block_sum(acc, red);
    if (threadIdx.x == 0) logjac[blockIdx.x] = tot;
}

// q2pos (Q): flat position i of the element that reads conditioner position q, or -1
__global__ void __launch_bounds__(kEncThreads)
coupling_x_to_z_bwd_kernel(int D, int Q, const float* __restrict__ z, const float* __restrict__ st,
                           const int32_t* __restrict__ pos2j, const int32_t* __restrict__ qmap,
                           const int32_t* __restrict__ q2pos, const float* __restrict__ g_z,
                           const float* __restrict__ g_lj, float* __restrict__ g_x, float* __restrict__ g_st) {
    const size_t xb = (size_t)blockIdx.x * D, sb = (size_t)blockIdx.x * 2 * Q;
    const float glj = g_lj ? g_lj[blockIdx.x] : 0.f;
    for (int i = threadIdx.x; i < D; i += kEncThreads) {
        const int j = __ldg(pos2j + i);
        float g = g_z ? g_z[xb + i] : 0.f;
        if (j >= 0) {
            const int q = qmap ? __ldg(qmap + j) : j;
            g *= expf(st[sb + Q + q]);
        }
        g_x[xb + i] = g;
    }
    for (int q = threadIdx.x; q < Q; q += kEncThreads) {
        const int i = __ldg(q2pos + q);
        float gt = 0.f, gs = 0.f;
        if (i >= 0) {
            const float g = g_z ? g_z[xb + i] : 0.f;
            gt = g * expf(st[sb + Q + q]);            // d z / d t = e^s
            gs = g * z[xb + i] + glj;                 // d z / d s = z, d logjac / d s = 1
        }
        g_st[sb + q] = gt;
        g_st[sb + Q + q] = gs;
    }
}

__global__ void __launch_bounds__(kEncThreads)
sumsq_diff_fwd_kernel(int n, const float* __restrict__ a, const float* __restrict__ c, float* __restrict__ out) {
    __shared__ float red[kEncThreads / 32];
    const size_t base = (size_t)blockIdx.x * n;
    float acc = 0.f;
    for (int i = threadIdx.x; i < n; i += kEncThreads) {
        const float d = a[base + i] - (c ? c[base + i] : 0.f);
        acc += d * d;
    }
    const float tot = block_sum(acc, red);
    if (threadIdx.x == 0) out[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(kEncThreads)
sumsq_diff_bwd_kernel(int n, const float* __restrict__ a, const float* __restrict__ c, const float* __restrict__ g,
                      float* __restrict__ g_a) {
    const size_t base = (size_t)blockIdx.x * n;
    const float gg = 2.f * g[blockIdx.x];
    for (int i = threadIdx.x; i < n; i += kEncThreads) g_a[base + i] = gg * (a[base + i] - (c ? c[base + i] : 0.f));
}

}  // namespace
}  // namespace cmf

using namespace cmf;

extern "C" int cmf_preproc_logit_fwd(int B, int n, float* x, const float* noise, float mult, float add,
                                     float* y, float* logjac, cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && n >= 0, "preproc_logit: negative size");
    CMF_REQUIRE(mult != 0.f, "preproc_logit: scalar multiplication by zero is not a bijection");
    if (B == 0) return 0;
    CMF_REQUIRE(x && y && logjac, "preproc_logit: null pointer");
    preproc_logit_kernel<<<(unsigned)B, kEncThreads, 0, (cudaStream_t)stream>>>(n, x, noise, mult, add,
                                                                               (float)n * logf(fabsf(mult)), y, logjac);
    CMF_CHECK_LAUNCH();
    return 0;
}

extern "C" int cmf_coupling_x_to_z_fwd(int B, int D, int Q, const float* x, const float* st, const int32_t* pos2j,
                                       const int32_t* qmap, float* z, float* logjac, cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && D >= 0 && Q >= 0, "coupling_x_to_z: negative size");
    if (B == 0) return 0;
    CMF_REQUIRE(x && st && pos2j && z && logjac, "coupling_x_to_z: null pointer");
    coupling_x_to_z_fwd_kernel<<<(unsigned)B, kEncThreads, 0, (cudaStream_t)stream>>>(D, Q, x, st, pos2j, qmap, z, logjac);
    CMF_CHECK_LAUNCH();
    return 0;
}

extern "C" int cmf_coupling_x_to_z_bwd(int B, int D, int Q, const float* z, const float* st, const int32_t* pos2j,
                                       const int32_t* qmap, const int32_t* q2pos, const float* g_z, const float* g_logjac,
                                       float* g_x, float* g_st, cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && D >= 0 && Q >= 0, "coupling_x_to_z_bwd: negative size");
    if (B == 0) return 0;
    CMF_REQUIRE(z && st && pos2j && q2pos && g_x && g_st, "coupling_x_to_z_bwd: null pointer");
    coupling_x_to_z_bwd_kernel<<<(unsigned)B, kEncThreads, 0, (cudaStream_t)stream>>>(D, Q, z, st, pos2j, qmap, q2pos, g_z,
                                                                                     g_logjac, g_x, g_st);
    CMF_CHECK_LAUNCH();
    return 0;
}

extern "C" int cmf_sumsq_diff_fwd(int B, int n, const float* a, const float* c, float* out, cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && n >= 0, "sumsq_diff: negative size");
    if (B == 0) return 0;
    CMF_REQUIRE(a && out, "sumsq_diff: null pointer");
    sumsq_diff_fwd_kernel<<<(unsigned)B, kEncThreads, 0, (cudaStream_t)stream>>>(n, a, c, out);
    CMF_CHECK_LAUNCH();
    return 0;
}

extern "C" int cmf_sumsq_diff_bwd(int B, int n, const float* a, const float* c, const float* g, float* g_a,
                                  cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && n >= 0, "sumsq_diff_bwd: negative size");
    if (B == 0) return 0;
    CMF_REQUIRE(a && g && g_a, "sumsq_diff_bwd: null pointer");
    sumsq_diff_bwd_kernel<<<(unsigned)B, kEncThreads, 0, (cudaStream_t)stream>>>(n, a, c, g, g_a);
    CMF_CHECK_LAUNCH();
    return 0;
}

// ---- per-channel sum over images and pixels: the bias gradient of a convolution -------------------------
// out[c] = sum_{n, p} x[n, c, p]   (autograd's `g_y.sum(dim = (0, 2, 3))` behind every biased Conv2d of
// cmf/models/components/networks.py:35-94; a generic torch reduction, 11 us per call in the step).
// One block per channel, fixed summation order (thread-strided partial sums, then a tree): reproducible.
namespace cmf {
namespace {
// stage 1: block (c, s) sums images n = s, s + S, ... of channel c; stage 2: the S partials in order
__global__ void __launch_bounds__(256) channel_sum_kernel(int N, int C, int HW, const float* __restrict__ x, float* __restrict__ partial) {
    __shared__ float r[256];
    const int c = blockIdx.x, S = gridDim.y, t = threadIdx.x;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    const int hw4 = (HW % 4 == 0) ? HW / 4 : 0;
    for (int n = blockIdx.y; n < N; n += S) {
        const float* row = x + ((long long)n * C + c) * HW;
        if (hw4) {
            const float4* row4 = reinterpret_cast<const float4*>(row);
            for (int i = t; i < hw4; i += 256) {
                const float4 v = __ldg(row4 + i);
                s0 += v.x; s1 += v.y; s2 += v.z; s3 += v.w;
            }
        } else {
            for (int i = t; i < HW; i += 256) s0 += __ldg(row + i);
        }
    }
    r[t] = (s0 + s1) + (s2 + s3);
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if (t < w) r[t] += r[t + w];
        __syncthreads();
    }
    if (t == 0) partial[c * S + blockIdx.y] = r[0];
}
__global__ void channel_sum_final_kernel(int C, int S, const float* __restrict__ partial, float* __restrict__ out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    float s = 0.f;
    for (int i = 0; i < S; ++i) s += partial[c * S + i];
    out[c] = s;
}
}  // namespace
}  // namespace cmf

extern "C" int cmf_channel_sum_splits(int N, int C) {
    if (N <= 0 || C <= 0) return 1;
    int s = (4 * cmf::kNumSMs + C - 1) / C;          // ~4 blocks per SM
    return s < 1 ? 1 : (s > N ? N : s);
}

extern "C" int cmf_channel_sum(int N, int C, int HW, const float* x, float* partial, float* out, cmf_stream_t stream) {
    using namespace cmf;
    CMF_REQUIRE(N >= 0 && C >= 1 && C <= 65535 && HW >= 1 && out && partial, "channel_sum: N = %d, C = %d, HW = %d", N, C, HW);
    CMF_REQUIRE(N == 0 || x, "channel_sum: null pointer");
    CMF_REQUIRE(N == 0 || HW % 4 != 0 || aligned16(x), "channel_sum: x must be 16-byte aligned");
    const int S = cmf_channel_sum_splits(N, C);
    channel_sum_kernel<<<dim3(C, S), 256, 0, (cudaStream_t)stream>>>(N, C, HW, x, partial);
    CMF_CHECK_LAUNCH();
    channel_sum_final_kernel<<<(C + 127) / 128, 128, 0, (cudaStream_t)stream>>>(C, S, partial, out);
    CMF_CHECK_LAUNCH();
    return 0;
}

// ---- ReLU of a primal activation together with its packed sign bits ------------------------------------
// relu_out[n, c, p] = max(x[n, c, p], 0) and bits[n, p] = the 64 sign bits of cmf_relu_mask_pack, in one pass
// over x (networks.py:62-79: every residual block applies the ReLU to the primal and, by jvp_layers.py:38-46,
// its 0/1 pattern to the tangents; the step ran a clamp kernel and a pack kernel over the same tensor).
namespace cmf {
namespace {
__global__ void relu_and_pack_kernel(long long total, int hw, const float* __restrict__ x, uint2* __restrict__ bits,
                                     float* __restrict__ relu_out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;      // = n * hw + p
    if (i >= total) return;
    const long long n = i / hw;
    const long long off = (n * 64) * hw + (i - n * hw);
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int e = 0; e < 32; ++e) {
        const float a = __ldg(x + off + (long long)e * hw), b = __ldg(x + off + (long long)(32 + e) * hw);
        lo |= (a > 0.f ? 1u : 0u) << e;
        hi |= (b > 0.f ? 1u : 0u) << e;
        relu_out[off + (long long)e * hw] = fmaxf(a, 0.f);
        relu_out[off + (long long)(32 + e) * hw] = fmaxf(b, 0.f);
    }
    bits[i] = make_uint2(lo, hi);
}
}  // namespace
}  // namespace cmf

extern "C" int cmf_relu_and_pack(int N, int HW, const float* x, uint32_t* bits, float* relu_out, cmf_stream_t stream) {
    using namespace cmf;
    CMF_REQUIRE(N >= 0 && HW >= 0, "relu_and_pack: negative size");
    if (N == 0 || HW == 0) return 0;
    CMF_REQUIRE(x && bits && relu_out, "relu_and_pack: null pointer");
    CMF_REQUIRE((reinterpret_cast<uintptr_t>(bits) & 7u) == 0, "relu_and_pack: bits must be 8-byte aligned");
    const long long total = (long long)N * HW;
    relu_and_pack_kernel<<<(unsigned)ceil_div(total, 256), 256, 0, (cudaStream_t)stream>>>(total, HW, x, reinterpret_cast<uint2*>(bits), relu_out);
    CMF_CHECK_LAUNCH();
    return 0;
}
