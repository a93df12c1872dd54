// Cosine-distance contraction on the 5th-generation tensor cores (sm_100a tcgen05 + TMEM).
//
//   out[ra[i]][j] = max(0, 1 - <A[ra[i]], B[rb[j]]> / (na[i] * nb[j]))      i < nA, j < nB
//
// -- embedding_distance of the reference (boxmot/utils/matching.py:230-321: cdist cosine of
// the track features against the detection features, clipped at 0) for one problem per
// blockIdx.y.  A and B rows are gathered through index lists (track slots / detection rows),
// fp32, K-major (a feature vector is contiguous).
//
// Precision: the operands are split x = hi + lo with hi = tf32(x) (rounded to nearest, low 13
// mantissa bits zero) and lo = tf32(x - hi); three kind::tf32 MMAs (hi.hi + hi.lo + lo.hi) accumulate in fp32 in
// TMEM, which reproduces the fp32 product to ~2^-20 relative ("3xTF32").  A single-pass tf32
// contraction (error ~1e-4 on the cosine) was not used: the distances feed an exact
// assignment whose result must equal the reference's.
//
// Structure per CTA (256 threads, one 128-row tile of A against up to 256 rows of B):
//   for each 32-element chunk of E:   all threads gather + split the chunk of the A and B
//     tiles into shared memory in the canonical K-major SWIZZLE_128B layout (8-row x 128-byte
//     atoms, 16-byte chunk index XOR row%8), fence.proxy.async, barrier; one thread issues the
//     4 x 3 tcgen05.mma (M=128, N=nB rounded up to 16, K=8) and commits to an mbarrier; all
//     threads wait on it before the buffers are overwritten.
//   epilogue: tcgen05.ld (32 lanes x 32 bit, 16 columns at a time; warp w reads the TMEM lane
//     quarter w % 4 and the column blocks of parity w / 4) -> cosine -> global.
// The operands are staged by ordinary loads instead of TMA because both tiles are GATHERED
// (rows are track slots / selected detections) and need the hi/lo split on the way in; at the
// tracker's shapes (A = hundreds of 2 KB rows per stream, read once) the kernel is bound by
// reading A from HBM, not by the tensor pipe.
#pragma once
#include "common.cuh"

namespace bm {

constexpr int TC_M = 128, TC_KC = 32, TC_NMAX = 256, TC_NT = 256;
constexpr uint32_t TC_TMEM_COLS = 256;

__host__ __device__ inline size_t cosine_tc_smem_bytes() {
    // fast path (cosine_tc_fast): 3 A stages + 2 B stages + 2 lo tiles = 106 KB; the plain path uses the first 96 KB
    return 1024 /*alignment slack*/ + (3 + 1) * (size_t)TC_M * 128 + (2 + 1) * (size_t)112 * 128 + 64;
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// start address >> 4 in [0,14), leading byte offset >> 4 in [16,30) (unused for this layout),
// stride byte offset >> 4 in [32,46) = 1024 B between 8-row groups, version 1 in [46,48),
// layout type 2 (SWIZZLE_128B) in [61,64).
__device__ __forceinline__ uint64_t tc_smem_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

// Instruction descriptor (cute::UMMA::InstrDescriptor): D = F32 (1 << 4), A = B = TF32
// (2 << 7, 2 << 10), both K-major, N >> 3 in [17,23), M >> 4 in [24,29).
__device__ __forceinline__ uint32_t tc_idesc(int n) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TC_M >> 4) << 24);
}

__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
        :: "r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ void tc_mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}\n" :: "r"(bar), "r"(parity) : "memory");
}

// Operand staging, split in two so that the global loads of chunk k+1 are in flight while the
// tensor core works on chunk k: tc_fetch gathers this thread's 16-byte pieces of a tile chunk
// into registers (piece q = tid + 128 j: row q / 8, 16-byte column q % 8), tc_store splits them
// into hi / lo and writes the swizzled buffers.
template <int NV>
__device__ __forceinline__ void tc_fetch(const float* const* rowp, int tile_rows, int k0, int kmax, bool vec, float4* v) {
#pragma unroll
    for (int j = 0; j < NV; j++) {
        const int q = threadIdx.x + TC_NT * j, c = q & 7;
        v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        const float* src = rowp[j];
        if ((q >> 3) < tile_rows && src) {
            src += k0 + c * 4;
            if (vec && k0 + c * 4 + 3 < kmax) v[j] = __ldg(reinterpret_cast<const float4*>(src));
            else {
                float t[4] = {0.f, 0.f, 0.f, 0.f};
                for (int e = 0; e < 4; e++) if (k0 + c * 4 + e < kmax) t[e] = src[e];
                v[j] = make_float4(t[0], t[1], t[2], t[3]);
            }
        }
    }
}

template <int NV>
__device__ __forceinline__ void tc_store(const float4* v, int tile_rows, unsigned char* hi, unsigned char* lo) {
#pragma unroll
    for (int j = 0; j < NV; j++) {
        const int q = threadIdx.x + TC_NT * j, r = q >> 3, c = q & 7;
        if (r >= tile_rows) continue;
        const float x[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
        uint32_t h[4], l[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            // round to nearest tf32 (add half an ulp of the 10-bit mantissa, clear the rest): the
            // split errors are then zero-mean instead of a systematic under-estimate
            h[e] = (__float_as_uint(x[e]) + 0x1000u) & 0xFFFFE000u;
            l[e] = (__float_as_uint(x[e] - __uint_as_float(h[e])) + 0x1000u) & 0xFFFFE000u;
        }
        const int off = r * 128 + ((c ^ (r & 7)) << 4);
        *reinterpret_cast<uint4*>(hi + off) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(lo + off) = make_uint4(l[0], l[1], l[2], l[3]);
    }
}

struct CosineProblem {
    const float* A; const int* ra; const float* na; int nA;    // rows ra[0..nA) of A, norms by position
    const float* B; const int* rb; const float* nb; int nB;
    int E;
    float* out; int out_ld;                                     // out[ra[i] * out_ld + j]
    int raw = 0;                                                // 1: out = <A, B> as it is (na / nb unused)
                                                                // 2: out[ra[i]] = min_j (1 - clip(cos, -1, 1)): nearest
                                                                //    neighbour over the B rows (StrongSORT's sample bank)
};

// ---------------------------------------------------------------------------------------------
// Fast path (E a multiple of 32, 16-byte aligned rows).  The plain loop below spends its time ISSUING
// INSTRUCTIONS (ncu: 35 % issue utilisation, ~525 warp instructions per 32-float chunk, two thirds of them
// the round-to-nearest hi / lo split and its address arithmetic), not waiting for memory.  Here:
//   * tcgen05 reads a tf32 operand from the upper 19 bits of an fp32 word, so the RAW fp32 tile is the `hi`
//     operand (hi = x truncated); cp.async drops the gathered 16-byte pieces straight into the canonical
//     K-major SWIZZLE_128B layout -- no register pass at all for hi;
//   * lo = x - trunc(x) (exact in fp32) costs one AND, one SUB and one ADD per element, from the landed tile;
//   * the A tile (bank rows: DRAM) has a ring of THREE stages, its copies are issued two chunks ahead; the B
//     tile (the frame's detections: L2 hits, every row tile of the stream re-reads them) a ring of two; one lo
//     tile each.  106 KB per CTA -> two CTAs per SM (tools/gather_bw.cu: residency beats ring depth), which is
//     why a column tile is 112 detections, not 128.
// Same three products (hi.hi + hi.lo + lo.hi), hi truncated instead of rounded, lo rounded; the cross terms
// accumulate in their own TMEM columns (see the MMA loop).  Error against scipy: tests/test_gpu_units.py.
constexpr int TC3_N = 112;
constexpr int TC3_RA = 3, TC3_RB = 2;
constexpr size_t TC3_AT = (size_t)TC_M * 128, TC3_BT = (size_t)TC3_N * 128;        // bytes of an A / B stage
constexpr size_t TC3_BYTES = TC3_RA * TC3_AT + TC3_RB * TC3_BT + TC3_AT + TC3_BT;   // + the two lo tiles
static_assert(TC3_AT % 1024 == 0 && TC3_BT % 1024 == 0, "SWIZZLE_128B tiles are 1024-byte aligned");

__device__ __forceinline__ void tc_cp16(uint32_t dst, const void* src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ uint4 tc_lo4(const uint4 x) {
    // + 0x1000: the tensor core drops the low 13 bits of lo as well; the half-ulp offset turns that truncation
    // into round-to-nearest
    uint4 l;
    l.x = __float_as_uint(__uint_as_float(x.x) - __uint_as_float(x.x & 0xFFFFE000u)) + 0x1000u;
    l.y = __float_as_uint(__uint_as_float(x.y) - __uint_as_float(x.y & 0xFFFFE000u)) + 0x1000u;
    l.z = __float_as_uint(__uint_as_float(x.z) - __uint_as_float(x.z & 0xFFFFE000u)) + 0x1000u;
    l.w = __float_as_uint(__uint_as_float(x.w) - __uint_as_float(x.w & 0xFFFFE000u)) + 0x1000u;
    return l;
}

__device__ __forceinline__ void cosine_tc_fast(const CosineProblem& pb, int r0, unsigned char* sm, uint32_t tmem,
                                               uint64_t* mbar) {
    const int tid = threadIdx.x, wid = tid >> 5, lane = tid & 31;
    unsigned char* hiA = sm;
    unsigned char* hiB = sm + TC3_RA * TC3_AT;
    unsigned char* loA = hiB + TC3_RB * TC3_BT;
    unsigned char* loB = loA + TC3_AT;
    const uint32_t bar = smem_u32(mbar);
    const int rows = min(TC_M, pb.nA - r0);
    const int nk = pb.E / TC_KC;
    constexpr int PA = TC_M * 8 / TC_NT, PB = (TC3_N * 8 + TC_NT - 1) / TC_NT;     // 16-byte pieces per thread and chunk
    float nn_min = CUDART_INF_F;
    uint32_t phase = 0;
    for (int c0 = 0; c0 < pb.nB; c0 += TC3_N) {
        const int cols = min(TC3_N, pb.nB - c0);
        const int npad = max(16, (cols + 15) & ~15);
        const uint32_t idesc = tc_idesc(npad);
        const float* srcA[PA];                                        // this thread's pieces: global source (nullptr = padding)
        const float* srcB[PB];
        int offA[PA], offB[PB];                                       // swizzled byte offset in a stage (B: -1 = beyond npad)
#pragma unroll
        for (int j = 0; j < PA; j++) {
            const int q = tid + TC_NT * j, r = q >> 3, c = q & 7;
            srcA[j] = r < rows ? pb.A + (size_t)pb.ra[r0 + r] * pb.E + c * 4 : nullptr;
            offA[j] = r * 128 + ((c ^ (r & 7)) << 4);
        }
#pragma unroll
        for (int j = 0; j < PB; j++) {
            const int q = tid + TC_NT * j, r = q >> 3, c = q & 7;
            srcB[j] = r < cols ? pb.B + (size_t)pb.rb[c0 + r] * pb.E + c * 4 : nullptr;
            offB[j] = r < npad ? r * 128 + ((c ^ (r & 7)) << 4) : -1;         // npad <= TC3_N covers q beyond the tile
        }
        // padding pieces are never copied: zero them in every hi stage once (lo follows from hi)
#pragma unroll
        for (int j = 0; j < PA; j++)
            if (!srcA[j])
                for (int i = 0; i < TC3_RA; i++) *reinterpret_cast<uint4*>(hiA + i * TC3_AT + offA[j]) = make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int j = 0; j < PB; j++)
            if (!srcB[j] && offB[j] >= 0)
                for (int i = 0; i < TC3_RB; i++) *reinterpret_cast<uint4*>(hiB + i * TC3_BT + offB[j]) = make_uint4(0, 0, 0, 0);
        // one commit group per call, in the order ... B(k) A(k + 1) | B(k + 1) A(k + 2) ...: "all but the newest
        // group" is then exactly "A(k) and B(k) have landed, A(k + 1) may still fly"
        auto issueA = [&](int k, int stage) {
            if (k < nk) {
                unsigned char* dst = hiA + (size_t)stage * TC3_AT;
#pragma unroll
                for (int j = 0; j < PA; j++)
                    if (srcA[j]) tc_cp16(smem_u32(dst + offA[j]), srcA[j] + (size_t)k * TC_KC);
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        auto issueB = [&](int k) {
            if (k < nk) {
                unsigned char* dst = hiB + (size_t)(k & 1) * TC3_BT;
#pragma unroll
                for (int j = 0; j < PB; j++)
                    if (srcB[j]) tc_cp16(smem_u32(dst + offB[j]), srcB[j] + (size_t)k * TC_KC);
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        issueA(0, 0);
        issueB(0);
        issueA(1, 1);
        int sa = 0;                                                   // k % TC3_RA
        for (int k = 0; k < nk; k++) {
            asm volatile("cp.async.wait_group 1;" ::: "memory");      // this thread's pieces of A(k) and B(k) have landed
            if (k > 0) { tc_mbar_wait(bar, phase); phase ^= 1u; }     // MMAs of chunk k - 1 done: its stages and lo are free
            issueB(k + 1);                                            // stage (k + 1) & 1 = that of chunk k - 1
            issueA(k + 2, sa == 0 ? 2 : sa - 1);                      // stage (k + 2) % 3 = that of chunk k - 1
            const unsigned char* ha = hiA + (size_t)sa * TC3_AT;
            const unsigned char* hb = hiB + (size_t)(k & 1) * TC3_BT;
#pragma unroll
            for (int j = 0; j < PA; j++)
                *reinterpret_cast<uint4*>(loA + offA[j]) = tc_lo4(*reinterpret_cast<const uint4*>(ha + offA[j]));
#pragma unroll
            for (int j = 0; j < PB; j++)
                if (offB[j] >= 0)
                    *reinterpret_cast<uint4*>(loB + offB[j]) = tc_lo4(*reinterpret_cast<const uint4*>(hb + offB[j]));
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t dah = tc_smem_desc(smem_u32(ha)), dal = tc_smem_desc(smem_u32(loA));
                const uint64_t dbh = tc_smem_desc(smem_u32(hb)), dbl = tc_smem_desc(smem_u32(loB));
#pragma unroll
                for (int kk = 0; kk < TC_KC / 8; kk++) {
                    const uint64_t adv = (uint64_t)((kk * 32) >> 4);
                    // hi.hi in columns [0, 128), the two cross terms in [128, 256): the tensor core TRUNCATES
                    // every fp32 accumulation (measured: ~3e-8 of bias per K = 8 step on a sum near 1), so the
                    // small terms get an accumulator of their own magnitude and the big one sees E / 8 additions
                    // instead of 3 E / 8
                    tc_mma_tf32(tmem, dah + adv, dbh + adv, idesc, (k > 0 || kk > 0) ? 1u : 0u);
                    tc_mma_tf32(tmem + 128u, dah + adv, dbl + adv, idesc, (k > 0 || kk > 0) ? 1u : 0u);
                    tc_mma_tf32(tmem + 128u, dal + adv, dbh + adv, idesc, 1u);
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                             :: "r"(bar) : "memory");
            }
            sa = sa == TC3_RA - 1 ? 0 : sa + 1;
        }
        tc_mbar_wait(bar, phase);
        phase ^= 1u;
        // ---- epilogue: warp w owns TMEM lanes 32 (w % 4) .. + 31 = tile rows
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int lq = wid & 3;
        const int r = lq * 32 + lane;
        const bool rin = r < rows;
        const float na = (rin && pb.raw != 1) ? pb.na[r0 + r] : 1.f;
        float* orow = (rin && pb.raw != 2) ? pb.out + (size_t)pb.ra[r0 + r] * pb.out_ld + c0 : nullptr;
        for (int cb = 16 * (wid >> 2); cb < npad; cb += 16 * (TC_NT / 128)) {
            uint32_t v[16];
            const uint32_t taddr = tmem + ((uint32_t)(lq * 32) << 16) + (uint32_t)cb;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                  "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                : "r"(taddr) : "memory");
            uint32_t w[16];
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]),
                  "=r"(w[8]), "=r"(w[9]), "=r"(w[10]), "=r"(w[11]), "=r"(w[12]), "=r"(w[13]), "=r"(w[14]), "=r"(w[15])
                : "r"(taddr + 128u) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int j = 0; j < 16; j++) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(w[j]));
            if (rin) {
                // fp32 from here on (the dot product is fp32-accurate; an fp64 divide per entry was 37 % of the
                // kernel's stall samples): 1 - dot / (|a| |b|), error ~1e-7 on top of the product's
                float o[16];
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    const int c = cb + j;
                    const float dot = __uint_as_float(v[j]);
                    if (pb.raw == 1) { o[j] = dot; continue; }
                    const float cs = c < cols ? dot / (na * pb.nb[c0 + c]) : 0.f;
                    if (pb.raw == 2) { if (c < cols) nn_min = fminf(nn_min, 1.f - fminf(1.f, fmaxf(-1.f, cs))); continue; }
                    o[j] = fmaxf(0.f, 1.f - cs);
                }
                if (pb.raw != 2) {
                    float* dst = orow + cb;
                    if (cb + 16 <= cols && ((uintptr_t)dst & 15) == 0) {
#pragma unroll
                        for (int j = 0; j < 16; j += 4)
                            *reinterpret_cast<float4*>(dst + j) = make_float4(o[j], o[j + 1], o[j + 2], o[j + 3]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; j++) if (cb + j < cols) dst[j] = o[j];
                    }
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
    }
    if (pb.raw == 2) {
        float* part = reinterpret_cast<float*>(loA);
        part[(wid >> 2) * TC_M + (wid & 3) * 32 + lane] = nn_min;
        __syncthreads();
        const int r = (wid & 3) * 32 + lane;
        if (wid < 4 && r < rows) pb.out[pb.ra[r0 + r]] = fminf(part[r], part[TC_M + r]);
    }
}

// grid = (row tiles of 128, problems); the caller provides problem(blockIdx.y).
template <typename ProblemFn>
__global__ void __launch_bounds__(TC_NT)
cosine_tc_kernel(ProblemFn problem) {
    extern __shared__ unsigned char tc_smem_raw[];
    const CosineProblem pb = problem((int)blockIdx.y);
    const int r0 = blockIdx.x * TC_M;
    if (r0 >= pb.nA || pb.nB <= 0) return;                      // uniform over the CTA
    const int tid = threadIdx.x, wid = tid >> 5, lane = tid & 31;
    unsigned char* sm = reinterpret_cast<unsigned char*>(((uintptr_t)tc_smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char* a_hi = sm;
    unsigned char* a_lo = a_hi + TC_M * 128;
    unsigned char* b_hi = a_lo + TC_M * 128;
    unsigned char* b_lo = b_hi + TC_NMAX * 128;
    uint64_t* mbar = reinterpret_cast<uint64_t*>(b_lo + TC_NMAX * 128);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 1);
    const uint32_t bar = smem_u32(mbar);
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (wid == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     :: "r"(smem_u32(tmem_slot)), "r"(TC_TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot;
    if ((pb.E % TC_KC) == 0 && (((uintptr_t)pb.A | (uintptr_t)pb.B) & 15) == 0) {          // uniform over the CTA
        __syncthreads();                              // everyone has read tmem_slot: the fast path reuses that memory
        uint64_t* mbar3 = reinterpret_cast<uint64_t*>(sm + TC3_BYTES);
        if (tid == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(smem_u32(mbar3)));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
        cosine_tc_fast(pb, r0, sm, tmem, mbar3);
        if (wid == 0)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(TC_TMEM_COLS) : "memory");
        return;
    }
    const int rows = min(TC_M, pb.nA - r0);
    uint32_t phase = 0;
    float nn_min = CUDART_INF_F;                                // raw == 2: running minimum of this thread's columns
    for (int c0 = 0; c0 < pb.nB; c0 += TC_NMAX) {
        const int cols = min(TC_NMAX, pb.nB - c0);
        const int npad = max(16, (cols + 15) & ~15);
        const uint32_t idesc = tc_idesc(npad);
        // row pointers of this thread's pieces (nullptr = zero padding)
        constexpr int NVA = TC_M * 8 / TC_NT, NVB = TC_NMAX * 8 / TC_NT;
        const float* pa[NVA];
        const float* pbp[NVB];
#pragma unroll
        for (int j = 0; j < NVA; j++) {
            const int r = (tid + TC_NT * j) >> 3;
            pa[j] = r < rows ? pb.A + (size_t)pb.ra[r0 + r] * pb.E : nullptr;
        }
#pragma unroll
        for (int j = 0; j < NVB; j++) {
            const int r = (tid + TC_NT * j) >> 3;
            pbp[j] = r < cols ? pb.B + (size_t)pb.rb[c0 + r] * pb.E : nullptr;
        }
        const bool vec = (pb.E & 3) == 0;
        float4 va[NVA], vb[NVB];
        tc_fetch<NVA>(pa, TC_M, 0, pb.E, vec, va);
        tc_fetch<NVB>(pbp, npad, 0, pb.E, vec, vb);
        for (int k0 = 0; k0 < pb.E; k0 += TC_KC) {
            tc_store<NVA>(va, TC_M, a_hi, a_lo);
            tc_store<NVB>(vb, npad, b_hi, b_lo);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t dah = tc_smem_desc(smem_u32(a_hi)), dal = tc_smem_desc(smem_u32(a_lo));
                const uint64_t dbh = tc_smem_desc(smem_u32(b_hi)), dbl = tc_smem_desc(smem_u32(b_lo));
#pragma unroll
                for (int kk = 0; kk < TC_KC / 8; kk++) {          // K = 8 tf32 = 32 bytes per instruction
                    const uint64_t adv = (uint64_t)((kk * 32) >> 4);
                    tc_mma_tf32(tmem, dah + adv, dbh + adv, idesc, (k0 > 0 || kk > 0) ? 1u : 0u);
                    tc_mma_tf32(tmem, dah + adv, dbl + adv, idesc, 1u);
                    tc_mma_tf32(tmem, dal + adv, dbh + adv, idesc, 1u);
                }
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                             :: "r"(bar) : "memory");
            }
            if (k0 + TC_KC < pb.E) {                              // next chunk's loads fly during the MMAs
                tc_fetch<NVA>(pa, TC_M, k0 + TC_KC, pb.E, vec, va);
                tc_fetch<NVB>(pbp, npad, k0 + TC_KC, pb.E, vec, vb);
            }
            tc_mbar_wait(bar, phase);
            phase ^= 1;
        }
        // ---- epilogue: warp w owns TMEM lanes 32w .. 32w+31 = tile rows
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int lq = wid & 3;                                   // TMEM lane quarter this warp may read
        const int r = lq * 32 + lane;
        const bool rin = r < rows;
        const float na = (rin && pb.raw != 1) ? pb.na[r0 + r] : 1.f;
        float* orow = (rin && pb.raw != 2) ? pb.out + (size_t)pb.ra[r0 + r] * pb.out_ld + c0 : nullptr;
        for (int cb = 16 * (wid >> 2); cb < npad; cb += 16 * (TC_NT / 128)) {
            uint32_t v[16];
            const uint32_t taddr = tmem + ((uint32_t)(lq * 32) << 16) + (uint32_t)cb;
            asm volatile(
                "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
                : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                  "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                : "r"(taddr) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            if (rin) {
#pragma unroll
                for (int j = 0; j < 16; j++) {
                    const int c = cb + j;
                    if (c < cols) {
                        if (pb.raw == 1) { orow[c] = __uint_as_float(v[j]); continue; }
                        const double cs = (double)__uint_as_float(v[j]) / ((double)na * (double)pb.nb[c0 + c]);
                        if (pb.raw == 2) { nn_min = fminf(nn_min, (float)(1.0 - fmin(1.0, fmax(-1.0, cs)))); continue; }
                        orow[c] = (float)fmax(0.0, 1.0 - cs);
                    }
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();       // every warp has drained its accumulator rows before the next column tile
    }
    if (pb.raw == 2) {
        // the two warp groups (column-block parity) hold partial minima of the same rows
        float* part = reinterpret_cast<float*>(a_hi);           // staging buffers are idle now
        part[(wid >> 2) * TC_M + (wid & 3) * 32 + lane] = nn_min;
        __syncthreads();
        const int r = (wid & 3) * 32 + lane;
        if (wid < 4 && r < rows) pb.out[pb.ra[r0 + r]] = fminf(part[r], part[TC_M + r]);
    }
    if (wid == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(tmem), "r"(TC_TMEM_COLS) : "memory");
    }
}

}  // namespace bm
