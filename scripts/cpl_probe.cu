// Stand-alone timing probe of the coupling adjoint (kernel (a) backward) at the scope-A shape.
// Compiles coupling.cu itself with whatever -D switches are under test:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 [-DCMF_CPL_BWD_...] -o cpl_probe scripts/cpl_probe.cu
// Prints the median time of one adjoint call (CUDA events around 10 calls on three rotating buffer sets, L2 flushed before).
#include "../cmf_b200/csrc/abi.cu"
#include "../cmf_b200/csrc/coupling.cu"

#include <algorithm>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

__global__ void fill_kernel(float* p, size_t n, unsigned seed) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        unsigned h = (unsigned)i * 2654435761u + seed;
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        p[i] = (float)(h & 0xffff) / 65536.f - 0.5f;
    }
}

__global__ void checksum_kernel(const unsigned* p, size_t n, unsigned long long* out) {
    unsigned long long acc = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        acc += (unsigned long long)p[i] * (unsigned long long)(2 * i + 1);
    atomicAdd(out, acc);
}

int main(int argc, char** argv) {
    const int B = argc > 1 ? atoi(argv[1]) : 512, d = argc > 2 ? atoi(argv[2]) : 64, M = argc > 3 ? atoi(argv[3]) : 392;
    const char* tag = argc > 4 ? argv[4] : "";
    const int NSET = 3;
    const size_t BM = (size_t)B * M, BdM = BM * d;
    struct Set { float *z, *V, *st, *ST, *gx, *gV, *gz, *gVin, *gst, *gST; } s[NSET];
    for (int i = 0; i < NSET; ++i) {
        float** ptrs[] = {&s[i].z, &s[i].V, &s[i].st, &s[i].ST, &s[i].gx, &s[i].gV, &s[i].gz, &s[i].gVin, &s[i].gst, &s[i].gST};
        const size_t sizes[] = {BM, BdM, 2 * BM, 2 * BdM, BM, BdM, BM, BdM, 2 * BM, 2 * BdM};
        for (int j = 0; j < 10; ++j) {
            CK(cudaMalloc(ptrs[j], sizes[j] * sizeof(float)));
            fill_kernel<<<1184, 256>>>(*ptrs[j], sizes[j], 17u * i + j);
        }
    }
    const size_t wsb = cmf_coupling_jvp_bwd_workspace(B, d, M, M);
    void* ws = nullptr;
    if (wsb) CK(cudaMalloc(&ws, wsb));
    char* flush = nullptr;
    CK(cudaMalloc(&flush, 256u << 20));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    std::vector<float> ts;
    for (int rep = 0; rep < 12; ++rep) {
        CK(cudaMemsetAsync(flush, rep, 256u << 20, 0));
        CK(cudaEventRecord(e0, 0));
        for (int c = 0; c < 10; ++c) {
            const Set& q = s[c % NSET];
            if (cmf_coupling_jvp_bwd(B, d, M, M, q.z, q.V, q.st, q.ST, nullptr, nullptr, q.gx, q.gV, q.gz, q.gVin, q.gst, q.gST, ws, wsb, nullptr)) {
                printf("launch failed: %s\n", cmf_last_error_string());
                return 1;
            }
        }
        CK(cudaEventRecord(e1, 0));
        CK(cudaDeviceSynchronize());
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep >= 3) ts.push_back(ms * 100.f);
    }
    std::sort(ts.begin(), ts.end());
    const double us = ts[ts.size() / 2], bytes = 24.0 * M * (1.0 + d) * B;
    // position-weighted checksum of the bit patterns of every output of set 0 (compare across builds / switches)
    unsigned long long* dsum = nullptr;
    unsigned long long hsum = 0;
    CK(cudaMalloc(&dsum, 8)); CK(cudaMemset(dsum, 0, 8));
    checksum_kernel<<<592, 256>>>((const unsigned*)s[0].gz, BM, dsum);
    checksum_kernel<<<592, 256>>>((const unsigned*)s[0].gst, 2 * BM, dsum);
    checksum_kernel<<<592, 256>>>((const unsigned*)s[0].gVin, BdM, dsum);
    checksum_kernel<<<592, 256>>>((const unsigned*)s[0].gST, 2 * BdM, dsum);
    CK(cudaMemcpy(&hsum, dsum, 8, cudaMemcpyDeviceToHost));
    printf("%-28s B=%d d=%d M=%d  %.2f us per call (min %.2f)  %.0f GB/s algorithmic  checksum %016llx\n", tag, B, d, M, us, ts[0],
           bytes / us * 1e-3, hsum);
    if (argc > 5) {   // calibration: the forward kernel and a device-to-device copy of the adjoint's bytes in the same harness
        for (int mode = 0; mode < 2; ++mode) {
            ts.clear();
            for (int rep = 0; rep < 12; ++rep) {
                CK(cudaMemsetAsync(flush, rep, 256u << 20, 0));
                CK(cudaEventRecord(e0, 0));
                for (int c = 0; c < 10; ++c) {
                    const Set& q = s[c % NSET];
                    if (mode == 0) cmf_coupling_jvp_fwd(B, d, M, M, q.z, q.V, q.st, q.ST, nullptr, q.gz, q.gVin, nullptr);
                    else CK(cudaMemcpyAsync(q.gST, q.ST, (size_t)(12.0 * M * (1.0 + d) * B), cudaMemcpyDeviceToDevice, 0));
                }
                CK(cudaEventRecord(e1, 0));
                CK(cudaDeviceSynchronize());
                float ms = 0.f;
                CK(cudaEventElapsedTime(&ms, e0, e1));
                if (rep >= 3) ts.push_back(ms * 100.f);
            }
            std::sort(ts.begin(), ts.end());
            const double u2 = ts[ts.size() / 2], by2 = mode == 0 ? 16.0 * M * (1.0 + d) * B : bytes;
            printf("  %-26s %.2f us per call  %.0f GB/s\n", mode == 0 ? "coupling forward" : "cudaMemcpy D2D, same bytes", u2, by2 / u2 * 1e-3);
        }
    }
    return 0;
}
