// Developer microbenchmark: what load bandwidth does the contraction's ACCESS PATTERN allow, with the tensor
// core, the hi/lo split and every barrier taken out?  Same grid as cosine_tc_kernel<SsBankProblem> (row tiles
// of 128 gathered bank rows x streams), same 16-byte cp.async pieces into a ring of 32 KB stages (128 A rows +
// 128 B rows x 128 bytes per chunk of 32 floats), nothing else.  Build: nvcc -O3 -gencode
// arch=compute_100a,code=sm_100a -o tools/gather_bw tools/gather_bw.cu ; run: tools/gather_bw [stages] [ctas/SM]
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <random>
#include <cuda_runtime.h>

constexpr int M = 128, N = 128, KC = 32, NT = 256;
constexpr size_t TILE = (size_t)(M + N) * 128;

__device__ __forceinline__ void cp16(unsigned dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
}

template <int R>
__global__ void __launch_bounds__(NT) gather_kernel(const float* A, const int* live, int nA, int Pcap, const float* B,
                                                     int nB, int Dcap, int E, int chunk_bytes, float* sink) {
    extern __shared__ __align__(1024) unsigned char sm[];
    const int s = blockIdx.y, r0 = blockIdx.x * M, tid = threadIdx.x;
    if (r0 >= nA) return;
    const int rows = min(M, nA - r0);
    const float* As = A + (size_t)s * Pcap * E;
    const float* Bs = B + (size_t)s * Dcap * E;
    const int* lv = live + (size_t)s * Pcap;
    const int ppr = chunk_bytes / 16;                    // 16-byte pieces per row and chunk
    const int npa = M * ppr / NT;                        // pieces per thread (A), same for B
    const int nk = E * 4 / chunk_bytes;
    const unsigned base = (unsigned)__cvta_generic_to_shared(sm);
    auto issue = [&](int k) {
        const unsigned dst = base + (unsigned)((k % R) * (size_t)(M + N) * chunk_bytes);
        for (int j = 0; j < npa; j++) {
            const int q = tid + NT * j, r = q / ppr, c = q % ppr;
            if (r < rows) cp16(dst + r * chunk_bytes + c * 16, As + (size_t)lv[r0 + r] * E + (size_t)k * (chunk_bytes / 4) + c * 4);
            if (r < nB) cp16(dst + M * chunk_bytes + r * chunk_bytes + c * 16, Bs + (size_t)r * E + (size_t)k * (chunk_bytes / 4) + c * 4);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int k = 0; k < R - 1 && k < nk; k++) issue(k);
    float acc = 0.f;
    for (int k = 0; k < nk; k++) {
        if (k + R - 1 < nk) issue(k + R - 1); else asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group %0;" :: "n"(R - 1) : "memory");
        acc += reinterpret_cast<const float*>(sm + (k % R) * (size_t)(M + N) * chunk_bytes)[tid];     // keep the stage "used"
        __syncthreads();
    }
    if (acc == 123.456f) sink[0] = acc;
}

int main(int argc, char** argv) {
    const int S = 256, Pcap = 8192, E = 768, Dcap = 128, nA = 3250, nB = 105;
    float *A, *B, *sink;
    int* live;
    cudaMalloc(&A, (size_t)S * Pcap * E * 4);
    cudaMalloc(&B, (size_t)S * Dcap * E * 4);
    cudaMalloc(&sink, 4);
    cudaMalloc(&live, (size_t)S * Pcap * 4);
    cudaMemset(A, 0, (size_t)S * Pcap * E * 4);
    cudaMemset(B, 0, (size_t)S * Dcap * E * 4);
    std::vector<int> h((size_t)S * Pcap);
    std::mt19937 rng(1);
    for (int s = 0; s < S; s++) {
        std::vector<int> p(Pcap);
        for (int i = 0; i < Pcap; i++) p[i] = i;
        std::shuffle(p.begin(), p.end(), rng);
        std::copy(p.begin(), p.end(), h.begin() + (size_t)s * Pcap);
    }
    cudaMemcpy(live, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    void* flush;
    cudaMalloc(&flush, 256u << 20);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const double gbytes = (double)S * nA * E * 4 / 1e9;
    auto run = [&](auto kern, int R, int chunk_bytes, const char* name) {
        const size_t smem = (size_t)R * (M + N) * chunk_bytes;
        if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) { printf("%s: smem %zu refused\n", name, smem); cudaGetLastError(); return; }
        float best = 1e9f;
        for (int it = 0; it < 4; it++) {
            cudaMemsetAsync(flush, it, 256u << 20);
            cudaEventRecord(e0);
            kern<<<dim3((nA + M - 1) / M, S), NT, smem>>>(A, live, nA, Pcap, B, nB, Dcap, E, chunk_bytes, sink);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            best = std::min(best, ms);
        }
        printf("%-28s stages %d chunk %4d B smem %6zu: %.3f ms  bank reads %.2f TB/s  (err %s)\n", name, R, chunk_bytes, smem, best,
               gbytes / best, cudaGetErrorString(cudaGetLastError()));
    };
    run(gather_kernel<2>, 2, 128, "2 stages (2-3 CTAs/SM)");
    run(gather_kernel<3>, 3, 128, "3 stages (2 CTAs/SM)");
    run(gather_kernel<4>, 4, 128, "4 stages (1 CTA/SM)");
    run(gather_kernel<6>, 6, 128, "6 stages (1 CTA/SM)");
    run(gather_kernel<2>, 2, 256, "2 stages, 256 B per row");
    run(gather_kernel<3>, 3, 256, "3 stages, 256 B per row");
    run(gather_kernel<2>, 2, 512, "2 stages, 512 B per row");
    return 0;
}
