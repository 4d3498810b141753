// Kernel (a): fused affine-coupling push of the primal and all d tangent columns, and its adjoint.
//
// HBM-bound streaming kernels.  Algorithmic traffic per launch (fp32):
//   fwd: reads z_m,s,t (3M) + V_m,S_dot,T_dot (3Md), writes x_m (M) + V_out (Md)  = 16 M (1+d) B bytes
//   bwd: reads g_V,V_m,S_dot (3Md) + O(M), writes g_Vin,g_S_dot,g_T_dot (3Md) + O(M) = 24 M (1+d) B bytes
// Thread mapping: one thread owns one 128-bit vector (4 consecutive compact elements j of one
// sample) and walks a chunk of tangent columns k, so exp(-s), z_m are computed once and reused
// from registers for every column; consecutive lanes touch consecutive 16-byte words of every
// stream (fully coalesced).  The mapped (checkerboard / permuted) variant does the same with
// 32-bit accesses, reading conditioner outputs through the int32 index map.
#include <mutex>

#include "common.cuh"

namespace cmf {

constexpr int kThreads = 256;
constexpr int kUnroll = 4;

// ------------------------------------------------------------------------------------------------
// forward, identity map, 128-bit path.  M == Q, M % 4 == 0, all pointers 16B aligned.
// grid = (ceil(B*M/4 / kThreads), KC)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
coupling_fwd_vec4(int B, int d, int M, int kchunk,
                  const float* __restrict__ z_m, const float* __restrict__ V_m,
                  const float* __restrict__ st, const float* __restrict__ ST,
                  float* __restrict__ x_m, float* __restrict__ V_out) {
    const int Mv = M >> 2;
    const long long t = (long long)blockIdx.x * kThreads + threadIdx.x;
    if (t >= (long long)B * Mv) return;
    const int b = (int)(t / Mv);
    const int j = (int)(t % Mv) << 2;
    const float4 z = ldg4(z_m + (size_t)b * M + j);
    const float4 s = ldg4(st + (size_t)b * 2 * M + M + j);
    float4 e;
    e.x = expf(-s.x); e.y = expf(-s.y); e.z = expf(-s.z); e.w = expf(-s.w);
    if (blockIdx.y == 0) {
        const float4 sh = ldg4(st + (size_t)b * 2 * M + j);
        float4 x;
        x.x = z.x * e.x - sh.x; x.y = z.y * e.y - sh.y; x.z = z.z * e.z - sh.z; x.w = z.w * e.w - sh.w;
        *reinterpret_cast<float4*>(x_m + (size_t)b * M + j) = x;
    }
    const int k0 = blockIdx.y * kchunk;
    const int k1 = min(d, k0 + kchunk);
    const float* vp = V_m + ((size_t)b * d + k0) * M + j;
    const float* tp = ST + ((size_t)b * d + k0) * 2 * M + j;
    float* op = V_out + ((size_t)b * d + k0) * M + j;
    int k = k0;
    for (; k + kUnroll <= k1; k += kUnroll) {
        float4 v[kUnroll], sd[kUnroll], td[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            v[u] = ldcs4(vp + (size_t)u * M);
            td[u] = ldcs4(tp + (size_t)u * 2 * M);
            sd[u] = ldcs4(tp + (size_t)u * 2 * M + M);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            float4 o;
            o.x = e.x * (v[u].x - z.x * sd[u].x) - td[u].x;
            o.y = e.y * (v[u].y - z.y * sd[u].y) - td[u].y;
            o.z = e.z * (v[u].z - z.z * sd[u].z) - td[u].z;
            o.w = e.w * (v[u].w - z.w * sd[u].w) - td[u].w;
            *reinterpret_cast<float4*>(op + (size_t)u * M) = o;
        }
        vp += (size_t)kUnroll * M; tp += (size_t)kUnroll * 2 * M; op += (size_t)kUnroll * M;
    }
    for (; k < k1; ++k) {
        const float4 v = ldcs4(vp), td = ldcs4(tp), sd = ldcs4(tp + M);
        float4 o;
        o.x = e.x * (v.x - z.x * sd.x) - td.x;
        o.y = e.y * (v.y - z.y * sd.y) - td.y;
        o.z = e.z * (v.z - z.z * sd.z) - td.z;
        o.w = e.w * (v.w - z.w * sd.w) - td.w;
        *reinterpret_cast<float4*>(op) = o;
        vp += M; tp += 2 * M; op += M;
    }
}

// ------------------------------------------------------------------------------------------------
// forward, general: optional index map, any M/Q, 32-bit accesses.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads)
coupling_fwd_map(int B, int d, int M, int Q, int kchunk,
                 const float* __restrict__ z_m, const float* __restrict__ V_m,
                 const float* __restrict__ st, const float* __restrict__ ST,
                 const int32_t* __restrict__ map,
                 float* __restrict__ x_m, float* __restrict__ V_out) {
    const long long t = (long long)blockIdx.x * kThreads + threadIdx.x;
    if (t >= (long long)B * M) return;
    const int b = (int)(t / M);
    const int j = (int)(t % M);
    const int q = map ? __ldg(map + j) : j;
    const float z = __ldg(z_m + (size_t)b * M + j);
    const float e = expf(-__ldg(st + (size_t)b * 2 * Q + Q + q));
    if (blockIdx.y == 0) x_m[(size_t)b * M + j] = z * e - __ldg(st + (size_t)b * 2 * Q + q);
    const int k0 = blockIdx.y * kchunk;
    const int k1 = min(d, k0 + kchunk);
    const float* vp = V_m + ((size_t)b * d + k0) * M + j;
    const float* tp = ST + ((size_t)b * d + k0) * 2 * Q + q;
    float* op = V_out + ((size_t)b * d + k0) * M + j;
    int k = k0;
    for (; k + kUnroll <= k1; k += kUnroll) {
        float v[kUnroll], sd[kUnroll], td[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            v[u] = __ldcs(vp + (size_t)u * M);
            td[u] = __ldg(tp + (size_t)u * 2 * Q);
            sd[u] = __ldg(tp + (size_t)u * 2 * Q + Q);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) op[(size_t)u * M] = e * (v[u] - z * sd[u]) - td[u];
        vp += (size_t)kUnroll * M; tp += (size_t)kUnroll * 2 * Q; op += (size_t)kUnroll * M;
    }
    for (; k < k1; ++k) {
        *op = e * (__ldcs(vp) - z * __ldg(tp + Q)) - __ldg(tp);
        vp += M; tp += 2 * Q; op += M;
    }
}

// ------------------------------------------------------------------------------------------------
// backward.  Per (b, j):  A = sum_k S_dot_k * g_k,  C = sum_k V_k * g_k   (g = g_V)
//   g_Vin_k   = e * g_k          g_S_dot_k = -e * z * g_k        g_T_dot_k = -g_k
//   g_z = e * (g_x - A)          g_s = -z e g_x - e (C - z A)    g_t = -g_x
// When the tangent range is split over blockIdx.y (KC > 1) the partial (A, C) go to the workspace
// [KC][2][B*M] and coupling_bwd_finalize reduces them in a fixed order (deterministic).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void finalize_store(float z, float e, float gx, float A, float C,
                                               float* g_z, float* g_t, float* g_s) {
    *g_z = e * (gx - A);
    *g_s = -z * e * gx - e * (C - z * A);
    *g_t = -gx;
}

// identity map, 128-bit path.  grid = (ceil(B*M/4 / kThreads), KC)
#ifndef CMF_CPL_BWD_MINB
#define CMF_CPL_BWD_MINB 2
#endif
#ifndef CMF_CPL_BWD_UNROLL
#define CMF_CPL_BWD_UNROLL 4
#endif
[[maybe_unused]] constexpr int kUnrollB = CMF_CPL_BWD_UNROLL;
// cache policy of the adjoint's streams (probe switches; scripts/cpl_probe.cu): loads 0 = ld.cs, 1 = ld.nc, 2 = ld.cg, 3 = ld.lu;
// stores 0 = st.cs, 1 = st (write-back, normal eviction), 2 = st.cg, 3 = st.wt
#ifndef CMF_CPL_BWD_LD
#define CMF_CPL_BWD_LD 0
#endif
#ifndef CMF_CPL_BWD_ST
#define CMF_CPL_BWD_ST 0
#endif
__device__ __forceinline__ float4 bwd_ld4(const float* p) {
    const float4* q = reinterpret_cast<const float4*>(p);
#if CMF_CPL_BWD_LD == 0
    return __ldcs(q);
#elif CMF_CPL_BWD_LD == 1
    return __ldg(q);
#elif CMF_CPL_BWD_LD == 2
    return __ldcg(q);
#else
    return __ldlu(q);
#endif
}
__device__ __forceinline__ void bwd_st4(float* p, float4 v) {
    float4* q = reinterpret_cast<float4*>(p);
#if defined(CMF_CPL_BWD_NOSTORE)
    if (v.x == 1.2345e-30f) *q = v;           // timing ablation only
#elif CMF_CPL_BWD_ST == 0
    __stcs(q, v);
#elif CMF_CPL_BWD_ST == 1
    *q = v;
#elif CMF_CPL_BWD_ST == 2
    __stcg(q, v);
#else
    __stwt(q, v);
#endif
}
// The adjoint's three input streams go through a per-thread ring in shared memory filled by 16-byte cp.async copies
// (LDGSTS: no registers held while a load is in flight): every thread copies only what it will read itself, so the ring
// needs cp.async.wait_group and no barrier.  kRingStages stages of kRingU tangent rows x 3 streams x 16 bytes x 256 threads;
// CMF_CPL_BWD_RING=0 (compile time) restores the register form (12 loads in flight per thread, then compute, then stores).
#ifndef CMF_CPL_BWD_RING
#define CMF_CPL_BWD_RING 1
#endif
#ifndef CMF_CPL_BWD_STAGES
#define CMF_CPL_BWD_STAGES 4
#endif
#ifndef CMF_CPL_BWD_U
#define CMF_CPL_BWD_U 2
#endif
constexpr int kRingStages = CMF_CPL_BWD_STAGES, kRingU = CMF_CPL_BWD_U;
constexpr int kRingStageVecs = 3 * kRingU * kThreads;                       // float4 per stage
constexpr size_t kBwdRingBytes = CMF_CPL_BWD_RING ? (size_t)kRingStages * kRingStageVecs * sizeof(float4) : 0;
static_assert((kRingStages & (kRingStages - 1)) == 0, "ring stages: a power of two");

__device__ __forceinline__ void cp_async16(uint32_t smem_dst, const float* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(kThreads, CMF_CPL_BWD_MINB)
coupling_bwd_vec4(int B, int d, int M, int kchunk,
                  const float* __restrict__ z_m, const float* __restrict__ V_m,
                  const float* __restrict__ st, const float* __restrict__ ST,
                  const float* __restrict__ g_x, const float* __restrict__ g_V,
                  float* __restrict__ g_z, float* __restrict__ g_Vin,
                  float* __restrict__ g_st, float* __restrict__ g_ST, float* __restrict__ ws, int S) {
    // S > 1 (2, 4 or 8): the tangent range is split over the WARPS of the CTA instead of blockIdx.y -- warp c * (8 / S) + w
    // owns chunk c of the CTA's column group w (32 consecutive 128-bit columns: 512 contiguous bytes per row, as in the
    // blockIdx.y form), and the S partial (A, C) of a column meet in shared memory in chunk order: no workspace, no second launch
    const int Mv = M >> 2;
    const long long tg = (long long)blockIdx.x * kThreads + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int wpc = (kThreads / 32) / S;         // column groups per CTA
    const int chunk = S > 1 ? warp / wpc : (int)blockIdx.y;
    const long long t_ = S > 1 ? ((long long)blockIdx.x * wpc + (warp - chunk * wpc)) * 32 + lane : tg;
    const bool live = t_ < (long long)B * Mv;
    if (!live && S == 1) return;
    const long long t = live ? t_ : 0;           // S > 1: idle threads still meet the barrier
    const int b = (int)(t / Mv);
    const int j = (int)(t % Mv) << 2;
    const float4 z = ldg4(z_m + (size_t)b * M + j);
    const float4 s = ldg4(st + (size_t)b * 2 * M + M + j);
    float4 e;
    e.x = expf(-s.x); e.y = expf(-s.y); e.z = expf(-s.z); e.w = expf(-s.w);
    float4 ez;
    ez.x = -e.x * z.x; ez.y = -e.y * z.y; ez.z = -e.z * z.z; ez.w = -e.w * z.w;
    float4 A = make_float4(0.f, 0.f, 0.f, 0.f), C = A;
    const int k0 = chunk * kchunk;
    const int k1 = live ? min(d, k0 + kchunk) : k0;
    const float* gp = g_V + ((size_t)b * d + k0) * M + j;
    const float* vp = V_m + ((size_t)b * d + k0) * M + j;
    const float* sp = ST + ((size_t)b * d + k0) * 2 * M + M + j;
    float* ovp = g_Vin + ((size_t)b * d + k0) * M + j;
    float* otp = g_ST + ((size_t)b * d + k0) * 2 * M + j;
    int k = k0;
#if CMF_CPL_BWD_RING
    {
        extern __shared__ __align__(16) float4 ring[];
        const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring) + threadIdx.x * 16u;
        const int n_it = k1 > k0 ? (k1 - k0) / kRingU : 0;
        auto issue = [&](int it) {
            const uint32_t dst = ring_s + (uint32_t)(it & (kRingStages - 1)) * (uint32_t)(kRingStageVecs * 16);
            const size_t row = (size_t)it * kRingU;
#pragma unroll
            for (int u = 0; u < kRingU; ++u) {
                cp_async16(dst + (uint32_t)((0 * kRingU + u) * kThreads * 16), gp + (row + u) * M);
                cp_async16(dst + (uint32_t)((1 * kRingU + u) * kThreads * 16), vp + (row + u) * M);
                cp_async16(dst + (uint32_t)((2 * kRingU + u) * kThreads * 16), sp + (row + u) * 2 * M);
            }
        };
#pragma unroll
        for (int it = 0; it < kRingStages - 1; ++it) {
            if (it < n_it) issue(it);
            cp_async_commit();
        }
#pragma unroll 1
        for (int it = 0; it < n_it; ++it) {
            if (it + kRingStages - 1 < n_it) issue(it + kRingStages - 1);
            cp_async_commit();
            cp_async_wait<kRingStages - 1>();          // stage `it` has landed (this thread's own copies: no barrier)
            const float4* src = ring + (size_t)(it & (kRingStages - 1)) * kRingStageVecs + threadIdx.x;
#pragma unroll
            for (int u = 0; u < kRingU; ++u) {
                const float4 g = src[(0 * kRingU + u) * kThreads], v = src[(1 * kRingU + u) * kThreads],
                             sd = src[(2 * kRingU + u) * kThreads];
                A.x += sd.x * g.x; A.y += sd.y * g.y; A.z += sd.z * g.z; A.w += sd.w * g.w;
                C.x += v.x * g.x;  C.y += v.y * g.y;  C.z += v.z * g.z;  C.w += v.w * g.w;
                bwd_st4(ovp + (size_t)u * M, make_float4(e.x * g.x, e.y * g.y, e.z * g.z, e.w * g.w));
                bwd_st4(otp + (size_t)u * 2 * M, make_float4(-g.x, -g.y, -g.z, -g.w));
                bwd_st4(otp + (size_t)u * 2 * M + M, make_float4(ez.x * g.x, ez.y * g.y, ez.z * g.z, ez.w * g.w));
            }
            ovp += (size_t)kRingU * M; otp += (size_t)kRingU * 2 * M;
        }
        cp_async_wait<0>();
        k = k0 + n_it * kRingU;
        gp += (size_t)n_it * kRingU * M; vp += (size_t)n_it * kRingU * M; sp += (size_t)n_it * kRingU * 2 * M;
    }
#else
    for (; k + kUnrollB <= k1; k += kUnrollB) {
        float4 g[kUnrollB], v[kUnrollB], sd[kUnrollB];
#pragma unroll
        for (int u = 0; u < kUnrollB; ++u) {
            g[u] = bwd_ld4(gp + (size_t)u * M);
            v[u] = bwd_ld4(vp + (size_t)u * M);
            sd[u] = bwd_ld4(sp + (size_t)u * 2 * M);
        }
#pragma unroll
        for (int u = 0; u < kUnrollB; ++u) {
            A.x += sd[u].x * g[u].x; A.y += sd[u].y * g[u].y; A.z += sd[u].z * g[u].z; A.w += sd[u].w * g[u].w;
            C.x += v[u].x * g[u].x;  C.y += v[u].y * g[u].y;  C.z += v[u].z * g[u].z;  C.w += v[u].w * g[u].w;
            bwd_st4(ovp + (size_t)u * M, make_float4(e.x * g[u].x, e.y * g[u].y, e.z * g[u].z, e.w * g[u].w));
            bwd_st4(otp + (size_t)u * 2 * M, make_float4(-g[u].x, -g[u].y, -g[u].z, -g[u].w));
            bwd_st4(otp + (size_t)u * 2 * M + M, make_float4(ez.x * g[u].x, ez.y * g[u].y, ez.z * g[u].z, ez.w * g[u].w));
        }
        gp += (size_t)kUnrollB * M; vp += (size_t)kUnrollB * M; sp += (size_t)kUnrollB * 2 * M;
        ovp += (size_t)kUnrollB * M; otp += (size_t)kUnrollB * 2 * M;
    }
#endif
#pragma unroll 1
    for (; k < k1; ++k) {
        const float4 g = bwd_ld4(gp), v = bwd_ld4(vp), sd = bwd_ld4(sp);
        A.x += sd.x * g.x; A.y += sd.y * g.y; A.z += sd.z * g.z; A.w += sd.w * g.w;
        C.x += v.x * g.x;  C.y += v.y * g.y;  C.z += v.z * g.z;  C.w += v.w * g.w;
        bwd_st4(ovp, make_float4(e.x * g.x, e.y * g.y, e.z * g.z, e.w * g.w));
        bwd_st4(otp, make_float4(-g.x, -g.y, -g.z, -g.w));
        bwd_st4(otp + M, make_float4(ez.x * g.x, ez.y * g.y, ez.z * g.z, ez.w * g.w));
        gp += M; vp += M; sp += 2 * M; ovp += M; otp += 2 * M;
    }
    if (S > 1) {
        // 0 + p0 + p1 + ... in chunk order, as coupling_bwd_finalize adds the workspace partials
        __shared__ float4 part[2][kThreads];
        part[0][threadIdx.x] = A;
        part[1][threadIdx.x] = C;
        __syncthreads();
        if (chunk != 0 || !live) return;
        A = make_float4(0.f, 0.f, 0.f, 0.f); C = A;
        for (int c = 0; c < S; ++c) {
            const float4 a = part[0][(c * wpc + warp) * 32 + lane], cc = part[1][(c * wpc + warp) * 32 + lane];
            A.x += a.x; A.y += a.y; A.z += a.z; A.w += a.w;
            C.x += cc.x; C.y += cc.y; C.z += cc.z; C.w += cc.w;
        }
    } else if (gridDim.y > 1) {
        const size_t BM = (size_t)B * M;
        float* w = ws + (size_t)blockIdx.y * 2 * BM + (size_t)b * M + j;
        *reinterpret_cast<float4*>(w) = A;
        *reinterpret_cast<float4*>(w + BM) = C;
        return;
    }
    {
        float4 gx = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g_x) gx = ldg4(g_x + (size_t)b * M + j);
        float4 gz, gt, gs;
        finalize_store(z.x, e.x, gx.x, A.x, C.x, &gz.x, &gt.x, &gs.x);
        finalize_store(z.y, e.y, gx.y, A.y, C.y, &gz.y, &gt.y, &gs.y);
        finalize_store(z.z, e.z, gx.z, A.z, C.z, &gz.z, &gt.z, &gs.z);
        finalize_store(z.w, e.w, gx.w, A.w, C.w, &gz.w, &gt.w, &gs.w);
        *reinterpret_cast<float4*>(g_z + (size_t)b * M + j) = gz;
        *reinterpret_cast<float4*>(g_st + (size_t)b * 2 * M + j) = gt;
        *reinterpret_cast<float4*>(g_st + (size_t)b * 2 * M + M + j) = gs;
    }
}

// general: thread per conditioner position (b, q); dense zero-filled writes of g_st / g_ST.
__global__ void __launch_bounds__(kThreads)
coupling_bwd_map(int B, int d, int M, int Q, int kchunk,
                 const float* __restrict__ z_m, const float* __restrict__ V_m,
                 const float* __restrict__ st, const float* __restrict__ ST,
                 const int32_t* __restrict__ inv,
                 const float* __restrict__ g_x, const float* __restrict__ g_V,
                 float* __restrict__ g_z, float* __restrict__ g_Vin,
                 float* __restrict__ g_st, float* __restrict__ g_ST, float* __restrict__ ws) {
    const long long t = (long long)blockIdx.x * kThreads + threadIdx.x;
    if (t >= (long long)B * Q) return;
    const int b = (int)(t / Q);
    const int q = (int)(t % Q);
    const int j = inv ? __ldg(inv + q) : q;
    const int k0 = blockIdx.y * kchunk;
    const int k1 = min(d, k0 + kchunk);
    float* otp = g_ST + ((size_t)b * d + k0) * 2 * Q + q;
    if (j < 0) {  // passthrough position of a checkerboard layer: conditioner output unused
        for (int k = k0; k < k1; ++k) { otp[0] = 0.f; otp[Q] = 0.f; otp += 2 * Q; }
        if (blockIdx.y == 0) { g_st[(size_t)b * 2 * Q + q] = 0.f; g_st[(size_t)b * 2 * Q + Q + q] = 0.f; }
        return;
    }
    const float z = __ldg(z_m + (size_t)b * M + j);
    const float e = expf(-__ldg(st + (size_t)b * 2 * Q + Q + q));
    const float ez = -e * z;
    float A = 0.f, C = 0.f;
    const float* gp = g_V + ((size_t)b * d + k0) * M + j;
    const float* vp = V_m + ((size_t)b * d + k0) * M + j;
    const float* sp = ST + ((size_t)b * d + k0) * 2 * Q + Q + q;
    float* ovp = g_Vin + ((size_t)b * d + k0) * M + j;
    int k = k0;
    for (; k + kUnroll <= k1; k += kUnroll) {
        float g[kUnroll], v[kUnroll], sd[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            g[u] = __ldcs(gp + (size_t)u * M);
            v[u] = __ldcs(vp + (size_t)u * M);
            sd[u] = __ldg(sp + (size_t)u * 2 * Q);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            A += sd[u] * g[u];
            C += v[u] * g[u];
            ovp[(size_t)u * M] = e * g[u];
            otp[(size_t)u * 2 * Q] = -g[u];
            otp[(size_t)u * 2 * Q + Q] = ez * g[u];
        }
        gp += (size_t)kUnroll * M; vp += (size_t)kUnroll * M; sp += (size_t)kUnroll * 2 * Q;
        ovp += (size_t)kUnroll * M; otp += (size_t)kUnroll * 2 * Q;
    }
    for (; k < k1; ++k) {
        const float g = __ldcs(gp), v = __ldcs(vp), sd = __ldg(sp);
        A += sd * g; C += v * g;
        *ovp = e * g; otp[0] = -g; otp[Q] = ez * g;
        gp += M; vp += M; sp += 2 * Q; ovp += M; otp += 2 * Q;
    }
    if (gridDim.y == 1) {
        const float gx = g_x ? __ldg(g_x + (size_t)b * M + j) : 0.f;
        finalize_store(z, e, gx, A, C, g_z + (size_t)b * M + j,
                       g_st + (size_t)b * 2 * Q + q, g_st + (size_t)b * 2 * Q + Q + q);
    } else {
        const size_t BM = (size_t)B * M;
        float* w = ws + (size_t)blockIdx.y * 2 * BM + (size_t)b * M + j;
        w[0] = A; w[BM] = C;
    }
}

// reduce the KC partials (fixed order) and emit the primal adjoints.  thread per (b, j).
__global__ void __launch_bounds__(kThreads)
coupling_bwd_finalize(int B, int M, int Q, int KC,
                      const float* __restrict__ z_m, const float* __restrict__ st,
                      const int32_t* __restrict__ map, const float* __restrict__ g_x,
                      const float* __restrict__ ws,
                      float* __restrict__ g_z, float* __restrict__ g_st) {
    const long long t = (long long)blockIdx.x * kThreads + threadIdx.x;
    const size_t BM = (size_t)B * M;
    if (t >= (long long)BM) return;
    const int b = (int)(t / M);
    const int j = (int)(t % M);
    const int q = map ? __ldg(map + j) : j;
    float A = 0.f, C = 0.f;
    for (int c = 0; c < KC; ++c) {
        A += ws[(size_t)c * 2 * BM + t];
        C += ws[(size_t)c * 2 * BM + BM + t];
    }
    const float z = __ldg(z_m + t);
    const float e = expf(-__ldg(st + (size_t)b * 2 * Q + Q + q));
    const float gx = g_x ? __ldg(g_x + t) : 0.f;
    finalize_store(z, e, gx, A, C, g_z + t, g_st + (size_t)b * 2 * Q + q, g_st + (size_t)b * 2 * Q + Q + q);
}

// How many tangent chunks to spread over blockIdx.y so the grid fills the machine (>= ~2 waves of
// 148 SMs x 2048 threads) even at the reference's default batch of 50.
static int pick_kchunks(long long threads_x, int d, long long resident_threads = 2048) {
    if (d <= 0) return 1;
    const long long want = 2LL * kNumSMs * resident_threads;
    long long kc = ceil_div(want, threads_x > 0 ? threads_x : 1);
    const long long kc_max = ceil_div(d, kUnroll);  // at least one unrolled batch per chunk
    if (kc > kc_max) kc = kc_max;
    if (kc < 1) kc = 1;
    return (int)kc;
}

static bool vec4_ok(int M, int Q, const void* map, std::initializer_list<const void*> ptrs) {
    if (map != nullptr || M != Q || (M & 3) != 0) return false;
    for (const void* p : ptrs)
        if (p != nullptr && !aligned16(p)) return false;
    return true;
}

}  // namespace cmf

using namespace cmf;

extern "C" int cmf_coupling_jvp_fwd(int B, int d, int M, int Q,
                                    const float* z_m, const float* V_m,
                                    const float* st, const float* ST, const int32_t* map,
                                    float* x_m, float* V_out, cmf_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    CMF_REQUIRE(B >= 0 && d >= 0 && M >= 0 && Q >= M, "coupling_fwd: bad sizes B=%d d=%d M=%d Q=%d", B, d, M, Q);
    CMF_REQUIRE(map != nullptr || M == Q, "coupling_fwd: identity map needs M == Q (M=%d Q=%d)", M, Q);
    if (B == 0 || M == 0) return 0;
    CMF_REQUIRE(z_m && st && x_m, "coupling_fwd: null primal pointer");
    CMF_REQUIRE(d == 0 || (V_m && ST && V_out), "coupling_fwd: null tangent pointer");
    if (vec4_ok(M, Q, map, {z_m, V_m, st, ST, x_m, V_out})) {
        const long long tx = (long long)B * (M >> 2);
        const int KC = pick_kchunks(tx, d);
        const int kchunk = d > 0 ? (int)ceil_div(ceil_div(d, KC), kUnroll) * kUnroll : 1;
        dim3 grid((unsigned)ceil_div(tx, kThreads), (unsigned)(d > 0 ? ceil_div(d, kchunk) : 1));
        coupling_fwd_vec4<<<grid, kThreads, 0, stream>>>(B, d, M, kchunk, z_m, V_m, st, ST, x_m, V_out);
    } else {
        const long long tx = (long long)B * M;
        const int KC = pick_kchunks(tx, d);
        const int kchunk = d > 0 ? (int)ceil_div(ceil_div(d, KC), kUnroll) * kUnroll : 1;
        dim3 grid((unsigned)ceil_div(tx, kThreads), (unsigned)(d > 0 ? ceil_div(d, kchunk) : 1));
        coupling_fwd_map<<<grid, kThreads, 0, stream>>>(B, d, M, Q, kchunk, z_m, V_m, st, ST, map, x_m, V_out);
    }
    CMF_CHECK_LAUNCH();
    return 0;
}

// The split over tangent chunks must be the same function of (B,d,M,Q) in the workspace query and
// in the launch; vectorisation (which changes threads_x) depends on pointers, so size for the worst case.
static int bwd_kchunks(int B, int d, int M, int Q, bool vec) {
    // the 128-bit adjoint holds 128 registers (two CTAs = 512 threads resident per SM): fewer, longer tangent chunks than
    // the forward's rule -- measured at B = 512, d = 64 (10 launches of scope A): 2048 -> 0.503 ms, 1024 / 768 / 512 -> 0.494-0.495,
    // 384 -> 0.508, 256 -> 0.525
    static const long long res = [] { const char* e = getenv("CMF_CPL_BWD_RESIDENT"); return e ? atoll(e) : 1024LL; }();
    return pick_kchunks(vec ? (long long)B * (M >> 2) : (long long)B * Q, d, vec ? res : 2048);
}

extern "C" size_t cmf_coupling_jvp_bwd_workspace(int B, int d, int M, int Q) {
    if (B <= 0 || M <= 0 || d <= 0) return 0;
    int kc = bwd_kchunks(B, d, M, Q, false);
    if (M == Q && (M & 3) == 0) kc = max(kc, bwd_kchunks(B, d, M, Q, true));
    return kc > 1 ? (size_t)kc * 2 * (size_t)B * M * sizeof(float) : 0;
}

extern "C" int cmf_coupling_jvp_bwd(int B, int d, int M, int Q,
                                    const float* z_m, const float* V_m,
                                    const float* st, const float* ST,
                                    const int32_t* map, const int32_t* inv,
                                    const float* g_x, const float* g_V,
                                    float* g_z, float* g_Vin, float* g_st, float* g_ST,
                                    void* workspace, size_t workspace_bytes, cmf_stream_t stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    CMF_REQUIRE(B >= 0 && d >= 0 && M >= 0 && Q >= M, "coupling_bwd: bad sizes B=%d d=%d M=%d Q=%d", B, d, M, Q);
    CMF_REQUIRE((map == nullptr) == (inv == nullptr), "coupling_bwd: map and inv must be given together");
    CMF_REQUIRE(map != nullptr || M == Q, "coupling_bwd: identity map needs M == Q (M=%d Q=%d)", M, Q);
    if (B == 0 || Q == 0) return 0;
    CMF_REQUIRE(z_m && st && g_z && g_st, "coupling_bwd: null primal pointer");
    CMF_REQUIRE(d == 0 || (V_m && ST && g_V && g_Vin && g_ST), "coupling_bwd: null tangent pointer");
    const bool vec = vec4_ok(M, Q, map, {z_m, V_m, st, ST, g_x, g_V, g_z, g_Vin, g_st, g_ST, workspace});
    const int KC = bwd_kchunks(B, d, M, Q, vec);
    const int kchunk = d > 0 ? (int)ceil_div(ceil_div(d, KC), kUnroll) * kUnroll : 1;
    const int ny = d > 0 ? (int)ceil_div(d, kchunk) : 1;
    if (ny > 1)
        CMF_REQUIRE(workspace && workspace_bytes >= (size_t)ny * 2 * (size_t)B * M * sizeof(float),
                    "coupling_bwd: workspace too small (%zu bytes, need %zu)", workspace_bytes,
                    (size_t)ny * 2 * (size_t)B * M * sizeof(float));
    float* ws = static_cast<float*>(workspace);
    if (vec && kBwdRingBytes + 8192 > 48 * 1024) {   // opt-in to the ring's dynamic shared memory (+ 8 KB static), once per device
        static DeviceOnce ring_attr;
        if (!ring_attr.done()) {
            cudaError_t e = cudaFuncSetAttribute(coupling_bwd_vec4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBwdRingBytes);
            if (e != cudaSuccess) { set_error("coupling_bwd attr: %s", cudaGetErrorString(e)); return (int)e; }
            ring_attr.mark();
        }
    }
    // 128-bit path, 2..8 tangent chunks: the chunks go to the warps of one CTA and meet in shared memory (one launch, no
    // workspace); CMF_CPL_BWD_WARPSPLIT=0 restores the blockIdx.y chunks + finalize launch
    static const bool warpsplit_off = [] { const char* e = getenv("CMF_CPL_BWD_WARPSPLIT"); return e && e[0] == '0'; }();
    static const bool skip_fin = [] { const char* e = getenv("CMF_CPL_BWD_SKIPFIN"); return e && e[0] == '1'; }();   // timing probe only
    // more than 8 chunks wanted (small batches): 8 warp chunks still win (B = 50, M = 392: 15.1 -> 10.8 us; B = 8: 10.6 -> 8.1)
    // unless the grid is far from filling the machine AND the per-thread loop is long (B = 1, d = 512, M = 1536: 19.2 -> 22.8)
    static const long long warpsplit_minfill = [] { const char* e = getenv("CMF_CPL_BWD_WARPSPLIT_MINFILL"); return e ? atoll(e) : 148LL * 128; }();
    const bool few_long = (long long)B * (M >> 2) * 8 < warpsplit_minfill && d > 256;
    if (vec && KC > 1 && !warpsplit_off && (KC <= 8 || !few_long)) {
        int S = 2;
        while (S < KC && S < 8) S <<= 1;
        const int kchunk_s = (int)ceil_div(ceil_div(d, S), kUnroll) * kUnroll;
        const unsigned nx = (unsigned)ceil_div((long long)B * (M >> 2), (kThreads / S));
        coupling_bwd_vec4<<<dim3(nx, 1), kThreads, kBwdRingBytes, stream>>>(B, d, M, kchunk_s, z_m, V_m, st, ST, g_x, g_V,
                                                               g_z, g_Vin, g_st, g_ST, nullptr, S);
        CMF_CHECK_LAUNCH();
        return 0;
    }
    if (vec) {
        dim3 grid((unsigned)ceil_div((long long)B * (M >> 2), kThreads), (unsigned)ny);
        coupling_bwd_vec4<<<grid, kThreads, kBwdRingBytes, stream>>>(B, d, M, kchunk, z_m, V_m, st, ST, g_x, g_V,
                                                         g_z, g_Vin, g_st, g_ST, ws, 1);
    } else {
        dim3 grid((unsigned)ceil_div((long long)B * Q, kThreads), (unsigned)ny);
        coupling_bwd_map<<<grid, kThreads, 0, stream>>>(B, d, M, Q, kchunk, z_m, V_m, st, ST, inv, g_x, g_V,
                                                        g_z, g_Vin, g_st, g_ST, ws);
    }
    CMF_CHECK_LAUNCH();
    if (ny > 1 && !skip_fin) {
        coupling_bwd_finalize<<<(unsigned)ceil_div((long long)B * M, kThreads), kThreads, 0, stream>>>(
            B, M, Q, ny, z_m, st, map, g_x, ws, g_z, g_st);
        CMF_CHECK_LAUNCH();
    }
    return 0;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of cmf_coupling_jvp_bwd (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA-graph capture,
// tight serving loops).  The tangent-chunk partials (A, C) live in a file-scope device array of a
// fixed, modest size that every GPU loading this library holds until exit.  All calls share it:
// calls that overlap on the GPU (two streams, two host threads) race on it.  Nothing in it has to
// start at zero -- every partial that coupling_bwd_finalize reads is written by the same call's
// first kernel -- so the contents left by the previous call are simply overwritten.
// ------------------------------------------------------------------------------------------------
namespace cmf {
constexpr size_t kStaticScratchBytes = CMF_COUPLING_BWD_STATIC_MAX_BYTES;
static __device__ __align__(256) float g_coupling_bwd_scratch[kStaticScratchBytes / sizeof(float)];

constexpr int kMaxDevices = 64;
static std::mutex g_scratch_mu;
static float* g_scratch_ptr[kMaxDevices] = {};

// Device address of the scratch array on the CURRENT device.  In host code the symbol's name is only
// a host-side stand-in; the address a kernel may use comes from cudaGetSymbolAddress, once per
// device, with that device current.
static int static_scratch(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        set_error("coupling_bwd_static: cudaGetDevice: %s", cudaGetErrorString(e));
        return (int)e;
    }
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("coupling_bwd_static: device ordinal %d out of range (max %d)", dev, kMaxDevices - 1);
        return -1;
    }
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (g_scratch_ptr[dev] == nullptr) {
        float* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_coupling_bwd_scratch);
        if (e != cudaSuccess || p == nullptr) {
            set_error("coupling_bwd_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return e != cudaSuccess ? (int)e : -1;
        }
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return 0;
}
}  // namespace cmf

extern "C" size_t cmf_coupling_jvp_bwd_static_max_bytes(void) { return kStaticScratchBytes; }

extern "C" int cmf_coupling_jvp_bwd_static(int B, int d, int M, int Q,
                                           const float* z_m, const float* V_m,
                                           const float* st, const float* ST,
                                           const int32_t* map, const int32_t* inv,
                                           const float* g_x, const float* g_V,
                                           float* g_z, float* g_Vin, float* g_st, float* g_ST,
                                           cmf_stream_t stream) {
    CMF_REQUIRE(B >= 0 && d >= 0 && M >= 0 && Q >= M, "coupling_bwd_static: bad sizes B=%d d=%d M=%d Q=%d", B, d, M, Q);
    const size_t need = cmf_coupling_jvp_bwd_workspace(B, d, M, Q);
    CMF_REQUIRE(need <= kStaticScratchBytes,
                "coupling_bwd_static: B=%d d=%d M=%d Q=%d needs %zu bytes of scratch, the static array holds %zu "
                "(use cmf_coupling_jvp_bwd with a caller-provided workspace)", B, d, M, Q, need, kStaticScratchBytes);
    float* ws = nullptr;
    if (need > 0) {
        const int rc = static_scratch(&ws);
        if (rc != 0) return rc;
    }
    return cmf_coupling_jvp_bwd(B, d, M, Q, z_m, V_m, st, ST, map, inv, g_x, g_V, g_z, g_Vin, g_st, g_ST,
                                ws, need > 0 ? kStaticScratchBytes : 0, stream);
}
