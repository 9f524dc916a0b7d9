// prefill.cuh — batched prompt ingestion (K2, K7 of SURVEY.md §2.3): the part of LM Studio's
// /api/v0/chat/completions work (nats_llm_studio.go:158-179) that really is a dense contraction.
//   * gemm_bf16_tcgen05: Y[M x N] (+)= X[M x K] . W[N x K]^T on the 5th-generation tensor cores: operands
//     staged by TMA (cp.async.bulk.tensor, 128-byte swizzle) through a 4-deep shared-memory ring, one elected
//     thread issues tcgen05.mma (kind::f16, bf16 x bf16 -> fp32) into a TMEM accumulator, tcgen05.commit
//     releases ring slots and signals the epilogue, which reads TMEM with tcgen05.ld.
//   * dequant_bf16_kernel: engine-layout quantised matrix -> bf16 rows (the GEMM's B operand).
//   * row-wise helpers: embedding gather, RMSNorm -> bf16, RoPE + KV-page write, SwiGLU, causal attention over
//     the paged fp16 KV cache.
// Numerics: bf16 operands, fp32 accumulation; the KV pages it writes are consumed by the decode kernel.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>

#include "decode.cuh"

namespace b200 {

constexpr int GEMM_BM = 128, GEMM_BN = 128, GEMM_BK = 64, GEMM_STAGES = 3;   // 3 stages = 97 KB: two CTAs per SM, one's epilogue overlaps the other's main loop
constexpr int GEMM_THREADS = 128;
constexpr uint32_t GEMM_STAGE_BYTES = (GEMM_BM + GEMM_BN) * GEMM_BK * 2;      // 32 KB
constexpr size_t GEMM_SMEM = GEMM_STAGES * GEMM_STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;

// ------------------------------------------------------------------------------------ tcgen05 / TMA wrappers
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tcgen05_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tcgen05_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem desc] . B[smem desc]^T, bf16 inputs, fp32 accumulate
__device__ __forceinline__ void tcgen05_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// shared-memory matrix descriptor: K-major tile, 128-byte swizzle, rows of 128 B, 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t umma_desc_sw128(const void* smem) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_u32(smem) & 0x3FFFF) >> 4);        // start address >> 4            bits [0,14)
    d |= (uint64_t)1 << 16;                                  // leading byte offset (unused for swizzled K-major)
    d |= (uint64_t)(1024 >> 4) << 32;                        // stride byte offset >> 4        bits [32,46)
    d |= (uint64_t)1 << 46;                                  // descriptor version (sm_100)
    d |= (uint64_t)2 << 61;                                  // layout: SWIZZLE_128B
    return d;
}
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int M, int N) {
    return (1u << 4)                 // D format: F32
         | (1u << 7)                 // A format: BF16
         | (1u << 10)                // B format: BF16
         | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);   // A, B K-major (bits 15,16 = 0)
}
__device__ __forceinline__ void tmem_ld_32x32b_x32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
          "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
          "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ------------------------------------------------------------------------------------ the GEMM
// C[M x N] fp32 row-major (ldc).  mode 0: C = acc * scale; mode 1: C += acc * scale (residual add).
// One 128x128 output tile per CTA; K is a multiple of 64; M, N edges are zero-filled by TMA and masked on store.
__global__ void __launch_bounds__(GEMM_THREADS, 2)
gemm_bf16_tcgen05(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, float* __restrict__ C,
                  int M, int N, int K, int ldc, int mode, float scale) {
    extern __shared__ uint8_t gemm_smem_raw[];
    uint8_t* base = (uint8_t*)(((uintptr_t)gemm_smem_raw + 1023) & ~(uintptr_t)1023);       // SWIZZLE_128B needs 1024-byte alignment
    uint8_t* sA = base; uint8_t* sB = base + GEMM_STAGES * GEMM_BM * GEMM_BK * 2;
    uint64_t* full = (uint64_t*)(base + GEMM_STAGES * GEMM_STAGE_BYTES);
    uint64_t* empty = full + GEMM_STAGES; uint64_t* tmem_full = empty + GEMM_STAGES; uint32_t* tmem_slot = (uint32_t*)(tmem_full + 1);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.y * GEMM_BM, n0 = blockIdx.x * GEMM_BN, nkb = K / GEMM_BK;
    if (tid == 0) {
        for (int i = 0; i < GEMM_STAGES; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        mbar_init(tmem_full, 1);
        fence_barrier_init();
    }
    if (warp == 0) tcgen05_alloc(tmem_slot, GEMM_BN);          // 128 fp32 columns x 128 lanes
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_acc = *tmem_slot;

    if (tid == 0) {                         // ===== TMA producer =====
        for (int kb = 0; kb < nkb; kb++) {
            const int s = kb % GEMM_STAGES; const uint32_t ph = (kb / GEMM_STAGES) & 1;
            mbar_wait(&empty[s], ph ^ 1);
            mbar_expect_tx(&full[s], GEMM_STAGE_BYTES);
            tma_load_2d(sA + s * GEMM_BM * GEMM_BK * 2, &tmap_a, kb * GEMM_BK, m0, &full[s]);
            tma_load_2d(sB + s * GEMM_BN * GEMM_BK * 2, &tmap_b, kb * GEMM_BK, n0, &full[s]);
        }
    } else if (tid == 32) {                 // ===== MMA issuer =====
        constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, GEMM_BN);
        for (int kb = 0; kb < nkb; kb++) {
            const int s = kb % GEMM_STAGES; const uint32_t ph = (kb / GEMM_STAGES) & 1;
            mbar_wait(&full[s], ph);
            tcgen05_fence_after();
            const uint64_t da = umma_desc_sw128(sA + s * GEMM_BM * GEMM_BK * 2), db = umma_desc_sw128(sB + s * GEMM_BN * GEMM_BK * 2);
#pragma unroll
            for (int k = 0; k < GEMM_BK / 16; k++)            // UMMA_K = 16 bf16 = 32 bytes: advance the start address by 2 (>>4 units)
                tcgen05_mma_f16(tmem_acc, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
            tcgen05_commit(&empty[s]);                         // frees the ring slot when these MMAs have read it
        }
        tcgen05_commit(tmem_full);                             // accumulator complete
    }
    // ===== epilogue: warp w owns TMEM lanes 32w..32w+31 = output rows m0 + 32w + lane =====
    mbar_wait(tmem_full, 0);
    __syncwarp();                                              // tcgen05.ld is warp-collective
    tcgen05_fence_after();
    const int row = m0 + warp * 32 + lane;
#pragma unroll 1
    for (int c = 0; c < GEMM_BN; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tmem_acc + ((uint32_t)(warp * 32) << 16) + c, r);
        if (row < M) {
            float* dst = C + (size_t)row * ldc + n0 + c;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                if (n0 + c + j + 3 < N) {
                    float4 v = make_float4(__uint_as_float(r[j]) * scale, __uint_as_float(r[j + 1]) * scale, __uint_as_float(r[j + 2]) * scale, __uint_as_float(r[j + 3]) * scale);
                    if (mode == 1) { const float4 o = *(const float4*)(dst + j); v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
                    *(float4*)(dst + j) = v;
                } else {
                    for (int e = 0; e < 4; e++) if (n0 + c + j + e < N) { float v = __uint_as_float(r[j + e]) * scale; if (mode == 1) v += dst[j + e]; dst[j + e] = v; }
                }
            }
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) tcgen05_dealloc(tmem_acc, GEMM_BN);
}

// ------------------------------------------------------------------------------------ the persistent GEMM
// Same product as gemm_bf16_tcgen05, organised for throughput: one CTA per SM walks the output tiles (m fastest, so the
// four row tiles of a 512-row chunk reuse a weight tile out of L2); 128 x BN tiles (BN = 256: the full-width UMMA shape,
// twice the arithmetic per shared-memory byte of 128 x 128); warp-specialised — one TMA thread, one MMA thread, four
// epilogue warps — with TWO accumulators in tensor memory, so the epilogue of tile i (tcgen05.ld, scale / residual add,
// stores) runs while the MMAs of tile i+1 are issued.
__device__ __forceinline__ uint32_t pack16x2(float a, float b, bool f16);      // defined with the row kernels below
constexpr int PG_THREADS = 192;                              // warp 0: TMA, warp 1: MMA + TMEM owner, warps 2..5: epilogue
template <int BN> struct PGemm {
    static constexpr int STAGES = BN == 256 ? 4 : BN >= 192 ? 5 : 6;
    static constexpr uint32_t ACC_STRIDE = BN <= 128 ? 128 : 256, TM_COLS = 2 * ACC_STRIDE;      // two accumulators; TMEM allocations are powers of two
    static constexpr uint32_t A_BYTES = GEMM_BM * GEMM_BK * 2, B_BYTES = BN * GEMM_BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr size_t SMEM = (size_t)STAGES * STAGE_BYTES + 1024 + 256;
};
template <int BN>
__global__ void __launch_bounds__(PG_THREADS, 1)
gemm_bf16_persistent(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, float* __restrict__ C,
                     int M, int N, int K, int ldc, int mode, float scale) {
    using G = PGemm<BN>;
    extern __shared__ uint8_t gemm_smem_raw[];
    uint8_t* base = (uint8_t*)(((uintptr_t)gemm_smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = base; uint8_t* sB = base + G::STAGES * G::A_BYTES;
    uint64_t* full = (uint64_t*)(base + (size_t)G::STAGES * G::STAGE_BYTES);
    uint64_t* empty = full + G::STAGES; uint64_t* acc_full = empty + G::STAGES; uint64_t* acc_empty = acc_full + 2; uint32_t* tmem_slot = (uint32_t*)(acc_empty + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m_tiles = (M + GEMM_BM - 1) / GEMM_BM, n_tiles = (N + BN - 1) / BN, tiles = m_tiles * n_tiles, nkb = K / GEMM_BK;
    if (tid == 0) {
        for (int i = 0; i < G::STAGES; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }      // four epilogue warps release an accumulator
        fence_barrier_init();
    }
    if (warp == 1) tcgen05_alloc(tmem_slot, G::TM_COLS);       // two accumulators of BN fp32 columns
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {                        // ===== TMA producer =====
        if (lane == 0) {
            uint32_t it = 0;
            for (int t = blockIdx.x; t < tiles; t += gridDim.x) {
                const int m0 = (t % m_tiles) * GEMM_BM, n0 = (t / m_tiles) * BN;
                for (int kb = 0; kb < nkb; kb++, it++) {
                    const int s = it % G::STAGES; const uint32_t ph = (it / G::STAGES) & 1;
                    mbar_wait(&empty[s], ph ^ 1);
                    mbar_expect_tx(&full[s], G::STAGE_BYTES);
                    tma_load_2d(sA + (size_t)s * G::A_BYTES, &tmap_a, kb * GEMM_BK, m0, &full[s]);
                    tma_load_2d(sB + (size_t)s * G::B_BYTES, &tmap_b, kb * GEMM_BK, n0, &full[s]);
                }
            }
        }
    } else if (warp == 1) {                 // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(GEMM_BM, BN);
            uint32_t it = 0, tl = 0;
            for (int t = blockIdx.x; t < tiles; t += gridDim.x, tl++) {
                const uint32_t a = tl & 1, aph = (tl >> 1) & 1;
                mbar_wait(&acc_empty[a], aph ^ 1);              // the epilogue has drained this accumulator (first use: passes)
                tcgen05_fence_after();
                const uint32_t tacc = tmem_base + a * G::ACC_STRIDE;
                for (int kb = 0; kb < nkb; kb++, it++) {
                    const int s = it % G::STAGES; const uint32_t ph = (it / G::STAGES) & 1;
                    mbar_wait(&full[s], ph);
                    tcgen05_fence_after();
                    const uint64_t da = umma_desc_sw128(sA + (size_t)s * G::A_BYTES), db = umma_desc_sw128(sB + (size_t)s * G::B_BYTES);
#pragma unroll
                    for (int k = 0; k < GEMM_BK / 16; k++) tcgen05_mma_f16(tacc, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
                    tcgen05_commit(&empty[s]);
                }
                tcgen05_commit(&acc_full[a]);
            }
        }
    } else {                                // ===== epilogue warps: warp w may touch TMEM lanes 32 (w % 4) .. +31 =====
        const int q = warp & 3;
        uint32_t tl = 0;
        for (int t = blockIdx.x; t < tiles; t += gridDim.x, tl++) {
            const int m0 = (t % m_tiles) * GEMM_BM, n0 = (t / m_tiles) * BN;
            const uint32_t a = tl & 1, aph = (tl >> 1) & 1;
            mbar_wait(&acc_full[a], aph);
            __syncwarp();
            tcgen05_fence_after();
            const int row = m0 + q * 32 + lane;
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(tmem_base + a * G::ACC_STRIDE + ((uint32_t)(q * 32) << 16) + c, r);
                if (row < M && mode == 2) {
                    // fused SwiGLU (gate/up rows are interleaved in the weight matrix, so columns 2i, 2i+1 are gate_i, up_i):
                    // ACT[row][i] = bf16(silu(gate) * up), the arithmetic of swiglu_rows_kernel; C is the bf16 output, ldc its row length
                    if (n0 + c < N) {                                  // N % 32 == 0 (checked by the host)
                        uint32_t o[8];
#pragma unroll
                        for (int k = 0; k < 8; k++) {
                            const float g0 = __uint_as_float(r[4 * k]), u0 = __uint_as_float(r[4 * k + 1]), g1 = __uint_as_float(r[4 * k + 2]), u1 = __uint_as_float(r[4 * k + 3]);
                            o[k] = pack16x2(g0 / (1.f + __expf(-g0)) * u0, g1 / (1.f + __expf(-g1)) * u1, false);
                        }
                        uint4* dst = (uint4*)((__nv_bfloat16*)C + (size_t)row * ldc + (n0 + c) / 2);
                        dst[0] = make_uint4(o[0], o[1], o[2], o[3]); dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
                    }
                } else if (row < M) {
                    float* dst = C + (size_t)row * ldc + n0 + c;
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        if (n0 + c + j + 3 < N) {
                            float4 v = make_float4(__uint_as_float(r[j]) * scale, __uint_as_float(r[j + 1]) * scale, __uint_as_float(r[j + 2]) * scale, __uint_as_float(r[j + 3]) * scale);
                            if (mode == 1) { const float4 o = *(const float4*)(dst + j); v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w; }
                            *(float4*)(dst + j) = v;
                        } else {
                            for (int e = 0; e < 4; e++) if (n0 + c + j + e < N) { float v = __uint_as_float(r[j + e]) * scale; if (mode == 1) v += dst[j + e]; dst[j + e] = v; }
                        }
                    }
                }
            }
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[a]);
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) tcgen05_dealloc(tmem_base, G::TM_COLS);
}

// which sequence slot and position each activation row belongs to: a prompt chunk (pos0 + m, one slot) or a
// decode batch (one row per sequence, explicit arrays)
struct RowMap {
    const int* pos; const int* slot; int pos0, slot0;
    __device__ __forceinline__ int p(int m) const { return pos ? pos[m] : pos0 + m; }
    __device__ __forceinline__ int s(int m) const { return slot ? slot[m] : slot0; }
};

// Activation rows that feed the dequant-in-GEMM kernel (qgemm.cuh) differ from the bf16 GEMM's in two ways, both
// selected by `perm` in the row kernels below: inside every aligned group of four K positions the elements are stored
// as (0, 2, 1, 3), and the 16-bit format is fp16 (clamped to the finite range) instead of bf16.
__host__ __device__ inline int kperm4(int k) { return (k & ~3) | ((k & 1) << 1) | ((k >> 1) & 1); }
__device__ __forceinline__ float clamp_h(float v) { return fminf(fmaxf(v, -65504.f), 65504.f); }
__device__ __forceinline__ uint32_t pack16x2(float a, float b, bool f16) {
    if (f16) { const __half2 h = __floats2half2_rn(clamp_h(a), clamp_h(b)); return *(const uint32_t*)&h; }
    const __nv_bfloat162 h = __floats2bfloat162_rn(a, b); return *(const uint32_t*)&h;
}
__device__ __forceinline__ unsigned short pack16(float a, bool f16) {
    return f16 ? __half_as_ushort(__float2half_rn(clamp_h(a))) : __bfloat16_as_ushort(__float2bfloat16(a));
}

// How a stream-K dequant-in-GEMM launch (qgemm.cuh) left its result: row tiles whose whole K range ran in one CTA are
// final in C; the others exist as fp32 partials of the contributing CTAs (consecutive CTA indices = ascending K).  The
// row kernel that consumes C sums them on the fly, in CTA order — no reduction pass, no atomics, deterministic.
// Programmatic dependent launch: the dequant-in-GEMM kernel that follows a row kernel is launched with
// programmaticStreamSerialization, so its CTAs may start (barrier init, TMEM allocation, weight ring fill — nothing that
// depends on this kernel) as soon as every CTA of the row kernel has passed this point; it waits (griddepcontrol.wait)
// before it touches activations or results.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

struct QgSplit {
    const float* ws;            // [cta][2 slots][64 batch rows][128 tile rows]; nullptr = nothing to gather
    int T, G, ks, n_seg, mode;  // total K-steps, CTAs, K-steps per tile, matrices in the launch, 1 = result is added to C
    int step0[3], col0[3];      // first global step / first output column of each matrix
    float scale;
};
constexpr int QG_WS_B = 64, QG_WS_N = 128;
// Stream-K arithmetic, one definition for the kernel, the gather and the self-test (32-bit: steps x CTAs < 2^31 is checked on
// the host).  CTA c owns the global K-steps [c*T/G, (c+1)*T/G); the CTA that owns step s is ceil((s+1)*G/T) - 1.
__host__ __device__ inline int qg_cta_first_step(int c, int T, int G) { return (int)((unsigned)c * (unsigned)T / (unsigned)G); }
__host__ __device__ inline int qg_cta_of_step(int s, int T, int G) { return (int)(((unsigned)(s + 1) * (unsigned)G + (unsigned)T - 1) / (unsigned)T) - 1; }
// v[0..N) = final values of C[b][col .. col+N) given `v` = what C holds now (N consecutive columns of one 4-aligned group)
template <int N>
__device__ __forceinline__ void qg_gather(const QgSplit& sp, int b, int col, float* v) {
    if (!sp.ws) return;
    int si = 0; while (si + 1 < sp.n_seg && col >= sp.col0[si + 1]) si++;
    const int nl = col - sp.col0[si], t0 = sp.step0[si] + (nl >> 7) * sp.ks;
    const int c_lo = qg_cta_of_step(t0, sp.T, sp.G), c_hi = qg_cta_of_step(t0 + sp.ks - 1, sp.T, sp.G);
    if (c_lo == c_hi) return;                                       // tile was not split: C is final
    float acc[N];
#pragma unroll
    for (int i = 0; i < N; i++) acc[i] = 0.f;
    for (int c0 = c_lo; c0 <= c_hi; c0 += 8) {                      // eight parts' loads in flight at a time, summed in CTA order
        float part[8][N];
#pragma unroll
        for (int j = 0; j < 8; j++) {
#pragma unroll
            for (int i = 0; i < N; i++) part[j][i] = 0.f;
            const int c = c0 + j;
            if (c <= c_hi) {
                const int cb = qg_cta_first_step(c, sp.T, sp.G);
                const float* src = sp.ws + ((size_t)(c * 2 + (cb >= t0 ? 0 : 1)) * QG_WS_B + b) * QG_WS_N + (nl & 127);
                if (N == 8) { const float4 p = __ldcg((const float4*)src), q = __ldcg((const float4*)src + 1); part[j][0] = p.x; part[j][1] = p.y; part[j][2 % N] = p.z; part[j][3 % N] = p.w; part[j][4 % N] = q.x; part[j][5 % N] = q.y; part[j][6 % N] = q.z; part[j][7 % N] = q.w; }
                else if (N == 4) { const float4 p = __ldcg((const float4*)src); part[j][0] = p.x; part[j][1] = p.y; part[j][2 % N] = p.z; part[j][3 % N] = p.w; }
                else if (N == 2) { const float2 p = __ldcg((const float2*)src); part[j][0] = p.x; part[j][1] = p.y; }
                else part[j][0] = __ldcg(src);
            }
        }
#pragma unroll
        for (int j = 0; j < 8; j++) if (c0 + j <= c_hi) {
#pragma unroll
            for (int i = 0; i < N; i++) acc[i] += part[j][i];
        }
    }
#pragma unroll
    for (int i = 0; i < N; i++) v[i] = sp.mode == 1 ? v[i] + acc[i] * sp.scale : acc[i] * sp.scale;
}

// ------------------------------------------------------------------------------------ row-wise kernels
// engine-layout quantised matrix -> bf16 [n_rows x K] (row r of the matrix at out + r*K).  One thread expands one run of
// 16 consecutive elements (16 packed bytes + the block's scales -> 32 bytes of bf16); float op order is the bit-exact
// dequantiser's (quant.cuh dequant_elem), the result is rounded to bf16 once.
__device__ __forceinline__ void dequant_run16(const uint8_t* __restrict__ w, int type, int n_grp, int row, int run, float* v) {
    const int gb = group_bytes(type), g = run >> 7, ug = run & 127;                       // 128 runs per 2048-element group
    const uint8_t* base = w + (((size_t)(row >> 1) * n_grp + g) * 2 + (row & 1)) * gb;
    if (type == T_Q8_0) {                                                                 // plane p = ug % 4 of lane ug / 4; block ug / 2
        const uint4 q = *(const uint4*)(base + (ug & 3) * 512 + (ug >> 2) * 16);
        const float d = h2f(*(const uint16_t*)(base + 2048 + (ug >> 1) * 2));
        const uint32_t x[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int i = 0; i < 16; i++) v[i] = __fmul_rn((float)(int8_t)(x[i >> 2] >> (8 * (i & 3))), d);
        return;
    }
    const int sb = ug >> 4, u = ug & 15;                                                  // super-block in group, run in super-block
    if (type == T_Q4_K || type == T_Q5_K) {
        const int bb = block_bytes(type), j = u >> 1, half = u & 1;                        // sub-block j (32 elements), its first/second 16
        const uint8_t* blk = base + sb * bb;
        const uint4 hd = *(const uint4*)blk;                                              // d, dmin, 12 engine-form scale bytes
        const uint32_t sw[3] = {hd.y, hd.z, hd.w}; const int j2 = j >> 1, bit = 24 * j2;  // 24-bit field of chunk j2
        const uint32_t f = bit % 32 == 0 ? sw[bit / 32] : (bit / 32 < 2 ? __funnelshift_r(sw[bit / 32], sw[bit / 32 + 1], bit % 32) : sw[2] >> (bit % 32));
        const int sc = (f >> (6 * (j & 1))) & 63, mn = (f >> (12 + 6 * (j & 1))) & 63;
        const float dl = __fmul_rn(h2f(hd.x & 0xFFFF), (float)sc), ml = __fmul_rn(h2f(hd.x >> 16), (float)mn);
        const uint4 q = *(const uint4*)(blk + (type == T_Q4_K ? 16 : 48) + 32 * j2 + 16 * half);
        uint4 hq = make_uint4(0, 0, 0, 0);
        if (type == T_Q5_K) hq = *(const uint4*)(blk + 16 + 16 * half);
        const uint32_t x[4] = {q.x, q.y, q.z, q.w}, hx[4] = {hq.x, hq.y, hq.z, hq.w};
#pragma unroll
        for (int i = 0; i < 16; i++) {
            int qv = (x[i >> 2] >> (8 * (i & 3) + 4 * (j & 1))) & 0xF;
            if (type == T_Q5_K) qv |= ((hx[i >> 2] >> (8 * (i & 3) + j)) & 1) << 4;
            v[i] = __fsub_rn(__fmul_rn(dl, (float)qv), ml);
        }
        return;
    }
    // Q6_K: run u -> 128-half n, 32-group grp, 16-run k; lane L = sb*4 + 2n + k holds its ql/qh bytes and scales
    const int n = u >> 3, grp = (u & 7) >> 1, k = u & 1, L = sb * 4 + 2 * n + k;
    const uint4 ql = *(const uint4*)(base + ((grp & 1) ? 512 : 0) + L * 16), qh = *(const uint4*)(base + 1024 + L * 16);
    const float d = h2f(*(const uint16_t*)(base + 1664 + sb * 2));
    const float dl = __fmul_rn(d, (float)(int8_t)base[1536 + L * 4 + grp]);
    const uint32_t x[4] = {ql.x, ql.y, ql.z, ql.w}, hx[4] = {qh.x, qh.y, qh.z, qh.w};
#pragma unroll
    for (int i = 0; i < 16; i++) {
        const int lo = (x[i >> 2] >> (8 * (i & 3) + (grp >= 2 ? 4 : 0))) & 0xF, hi = (hx[i >> 2] >> (8 * (i & 3) + 2 * grp)) & 3;
        v[i] = __fmul_rn(dl, (float)((lo | (hi << 4)) - 32));
    }
}
__global__ void dequant_bf16_kernel(const uint8_t* __restrict__ w, int type, long long n_rows, int K, __nv_bfloat16* __restrict__ out) {
    const int ng = n_groups(K), rpr = K / 16;                                             // runs per row
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_rows * rpr; i += (long long)gridDim.x * blockDim.x) {
        const int row = (int)(i / rpr), run = (int)(i % rpr);
        float v[16]; dequant_run16(w, type, ng, row, run, v);
        uint32_t o[8];
#pragma unroll
        for (int j = 0; j < 8; j++) o[j] = pack16x2(v[2 * j], v[2 * j + 1], false);
        uint4* dst = (uint4*)(out + (size_t)row * K + run * 16);
        dst[0] = make_uint4(o[0], o[1], o[2], o[3]); dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
    }
}

// H[m][:] = dequant(token_embd[tok[m]]) * emb_scale
__global__ void embed_rows_kernel(const uint8_t* __restrict__ embd, int type, int D, const int* __restrict__ toks, int M, float emb_scale, float* __restrict__ H) {
    const int ng = n_groups(D);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)M * D; i += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(i / D), e = (int)(i % D);
        float v = embed_elem(embd, type, ng, toks[m], e);
        H[i] = emb_scale != 0.f ? v * emb_scale : v;
    }
}

// X[m][:] = 16-bit(rmsnorm(H[m]) * w); one block per row.  With a split producer (sp.ws) H[m] is first completed from
// the partials of the residual GEMM that precedes the norm and written back.
__global__ void __launch_bounds__(1024) rmsnorm_rows_kernel(float* __restrict__ H, const float* __restrict__ w, int D, int M, float eps, __nv_bfloat16* __restrict__ X, bool perm,
                                                           const __grid_constant__ QgSplit sp) {
    __shared__ float red[32];
    pdl_launch_dependents(); pdl_wait();
    const int m = blockIdx.x, tid = threadIdx.x, nth = blockDim.x;
    float* h = H + (size_t)m * D; float ss = 0.f;
    for (int i = tid * 4; i < D; i += nth * 4) {
        float4 v = *(const float4*)(h + i);
        if (sp.ws) { float t[4] = {v.x, v.y, v.z, v.w}; qg_gather<4>(sp, m, i, t); v = make_float4(t[0], t[1], t[2], t[3]); *(float4*)(h + i) = v; }
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    ss = warp_sum(ss);
    if ((tid & 31) == 0) red[tid >> 5] = ss;
    __syncthreads();
    ss = 0.f;
    for (int i = 0; i < (nth >> 5); i++) ss += red[i];              // same order in every thread
    const float rs = rsqrtf(ss / (float)D + eps);
    for (int i = tid * 4; i < D; i += nth * 4) {                     // each thread re-reads exactly what it wrote above
        const float4 v = *(const float4*)(h + i), g = *(const float4*)(w + i);
        const float a = v.x * rs * g.x, b = v.y * rs * g.y, c = v.z * rs * g.z, d = v.w * rs * g.w;
        *(uint2*)(X + (size_t)m * D + i) = perm ? make_uint2(pack16x2(a, c, true), pack16x2(b, d, true)) : make_uint2(pack16x2(a, b, false), pack16x2(c, d, false));
    }
}

// QKV[m][QD + 2*KVD] fp32 -> RoPE(q) as bf16 Q[m][QD]; RoPE(k), v -> fp16 KV pages at position pos0 + m
__global__ void rope_kv_rows_kernel(const float* __restrict__ QKV, int M, RowMap rm, int layer, int n_head, int n_head_kv, int HD, int n_rot,
                                    const float* __restrict__ inv_freq, KvLayout kv, __nv_bfloat16* __restrict__ Q, const __grid_constant__ QgSplit sp, bool q_f16) {
    pdl_launch_dependents(); pdl_wait();
    const int QD = n_head * HD, KVD = n_head_kv * HD, W = QD + 2 * KVD;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)M * (W / 2); i += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(i / (W / 2)), c = (int)(i % (W / 2)) * 2, pos = rm.p(m), slot = rm.s(m);
        float a = QKV[(size_t)m * W + c], b = QKV[(size_t)m * W + c + 1];
        if (sp.ws) { float t[2] = {a, b}; qg_gather<2>(sp, m, c, t); a = t[0]; b = t[1]; }
        if (c < QD + KVD) {                               // q or k: rotate the adjacent pair
            const int d = c % HD;
            if (d < n_rot) { float sn, cs; sincosf((float)pos * inv_freq[d >> 1], &sn, &cs); const float x = a, y = b; a = x * cs - y * sn; b = x * sn + y * cs; }
        }
        if (c < QD) { if (q_f16) *(__half2*)((__half*)Q + (size_t)m * QD + c) = __floats2half2_rn(a, b); else *(__nv_bfloat162*)(Q + (size_t)m * QD + c) = __floats2bfloat162_rn(a, b); }      // fp16 for the tensor-core prompt attention (same type as the KV cache)
        else if (c < QD + KVD) { const int r = c - QD; *(__half2*)(kv.at_slot(slot, layer, 0, r / HD, pos) + r % HD) = __floats2half2_rn(a, b); }
        else { const int r = c - QD - KVD; *(__half2*)(kv.at_slot(slot, layer, 1, r / HD, pos) + r % HD) = __floats2half2_rn(a, b); }
    }
}

// causal attention for the M new rows (positions pos0..pos0+M-1) over the paged KV cache; one warp per (row, head)
template <int VEC>
__global__ void attn_prefill_kernel(const __nv_bfloat16* __restrict__ Q, int M, RowMap rm, int layer, int n_head, int n_head_kv, float kq_scale,
                                    KvLayout kv, __nv_bfloat16* __restrict__ O, bool perm) {
    constexpr int HD = 32 * VEC, NB = 8;
    const int gw = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (gw >= M * n_head) return;
    const int m = gw / n_head, head = gw % n_head, kvh = head / (n_head / n_head_kv), n_ctx = rm.p(m) + 1, slot = rm.s(m);
    float qv[VEC], acc[VEC];
#pragma unroll
    for (int i = 0; i < VEC; i++) { qv[i] = __bfloat162float(Q[(size_t)m * n_head * HD + head * HD + lane * VEC + i]) * kq_scale; acc[i] = 0.f; }
    float mx = -INFINITY, l = 0.f;
    for (int pb = 0; pb < n_ctx; pb += NB) {
        float kf[NB][VEC], vf[NB][VEC];
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const int p = pb + b;
            if (p < n_ctx) {
                const __half* kp = kv.at_slot(slot, layer, 0, kvh, p) + lane * VEC; const __half* vp = kv.at_slot(slot, layer, 1, kvh, p) + lane * VEC;
#pragma unroll
                for (int i = 0; i < VEC; i += 2) {
                    const float2 a = __half22float2(*(const __half2*)(kp + i)), c2 = __half22float2(*(const __half2*)(vp + i));
                    kf[b][i] = a.x; kf[b][i + 1] = a.y; vf[b][i] = c2.x; vf[b][i + 1] = c2.y;
                }
            }
        }
#pragma unroll
        for (int b = 0; b < NB; b++) {
            if (pb + b >= n_ctx) break;
            float s = 0.f;
#pragma unroll
            for (int i = 0; i < VEC; i++) s += qv[i] * kf[b][i];
            s = warp_sum(s);
            const float mn = fmaxf(mx, s), corr = __expf(mx - mn), pe = __expf(s - mn);
            l = l * corr + pe;
#pragma unroll
            for (int i = 0; i < VEC; i++) acc[i] = acc[i] * corr + pe * vf[b][i];
            mx = mn;
        }
    }
    const float inv = 1.f / l;
#pragma unroll
    for (int i = 0; i < VEC; i++) { const int e = lane * VEC + i; ((unsigned short*)O)[(size_t)m * n_head * HD + head * HD + (perm ? kperm4(e) : e)] = pack16(acc[i] * inv, perm); }
}

// Attention for M rows over the paged fp16 KV cache, one block per (row, kv head): the G query heads that share the kv
// head are served from one pass over K and V.  Each of the 8 warps takes every eighth 32-position chunk; scores are
// computed one position per lane (no shuffles in the dot products), probabilities are broadcast by shuffle for the
// P.V accumulation (one output-dim slice per lane, eight V rows in flight), and the warps' online-softmax states are
// merged in shared memory.
constexpr int ATTN_WARPS = 4;
__device__ __forceinline__ float2 h2_to_f2(uint32_t u) { return __half22float2(*reinterpret_cast<const __half2*>(&u)); }
template <int HD, int G>
__global__ void __launch_bounds__(ATTN_WARPS * 32, 4) attn_rows_kernel(const __nv_bfloat16* __restrict__ Q, int M, RowMap rm, int layer, int n_head, int n_head_kv,
                                                                   float kq_scale, KvLayout kv, __nv_bfloat16* __restrict__ O, bool perm) {
    constexpr int DPL = HD / 32, NW = ATTN_WARPS;
    __shared__ __align__(16) float sq[G][HD];
    __shared__ float sm_m[NW][G], sm_l[NW][G];
    __shared__ float sm_acc[NW][G][HD];
    pdl_launch_dependents(); pdl_wait();
    const int m = blockIdx.x / n_head_kv, kvh = blockIdx.x % n_head_kv, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_ctx = rm.p(m) + 1, slot = rm.s(m);
    for (int i = tid; i < G * HD; i += NW * 32) { const int g = i / HD, d = i % HD; sq[g][d] = __bfloat162float(Q[(size_t)m * n_head * HD + (kvh * G + g) * HD + d]) * kq_scale; }
    __syncthreads();
    float mx[G], l[G], acc[G][DPL];
#pragma unroll
    for (int g = 0; g < G; g++) { mx[g] = -INFINITY; l[g] = 0.f;
#pragma unroll
        for (int i = 0; i < DPL; i++) acc[g][i] = 0.f; }
    for (int p0 = warp * 32; p0 < n_ctx; p0 += NW * 32) {
        const int p = p0 + lane; const bool ok = p < n_ctx;
        float s[G];
#pragma unroll
        for (int g = 0; g < G; g++) s[g] = 0.f;
        {
            const __half* kp = kv.at_slot(slot, layer, 0, kvh, ok ? p : n_ctx - 1);      // clamped: masked below
            constexpr int KB = HD >= 128 ? 8 : 4;                                   // K pieces in flight per lane; at least two trips, or the
#pragma unroll 1                                                                   // compiler hoists every q load of the chunk and spills
            for (int j0 = 0; j0 < HD / 8; j0 += KB) {
                uint4 kk[KB];
#pragma unroll
                for (int j = 0; j < KB; j++) kk[j] = *(const uint4*)(kp + 8 * (j0 + j));
#pragma unroll
                for (int j = 0; j < KB; j++) {
                    const float2 k0 = h2_to_f2(kk[j].x), k1 = h2_to_f2(kk[j].y), k2 = h2_to_f2(kk[j].z), k3 = h2_to_f2(kk[j].w);
#pragma unroll
                    for (int g = 0; g < G; g++) {
                        const float4 qa = *(const float4*)&sq[g][8 * (j0 + j)], qb = *(const float4*)&sq[g][8 * (j0 + j) + 4];
                        s[g] += qa.x * k0.x + qa.y * k0.y + qa.z * k1.x + qa.w * k1.y + qb.x * k2.x + qb.y * k2.y + qb.z * k3.x + qb.w * k3.y;
                    }
                }
            }
        }
        float pe[G];
#pragma unroll
        for (int g = 0; g < G; g++) {
            const float sg = ok ? s[g] : -INFINITY;
            float cm = sg;
#pragma unroll
            for (int o = 16; o; o >>= 1) cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, o));
            const float mn = fmaxf(mx[g], cm), corr = __expf(mx[g] - mn);
            pe[g] = ok ? __expf(sg - mn) : 0.f;
            l[g] = l[g] * corr + warp_sum(pe[g]);
#pragma unroll
            for (int i = 0; i < DPL; i++) acc[g][i] *= corr;
            mx[g] = mn;
        }
        // P.V: masked positions carry probability 0, so their (clamped, valid) V rows may be read unconditionally
        const int last = n_ctx - 1 - p0;                                           // >= 0
        const __half* vb0 = kv.at_slot(slot, layer, 1, kvh, p0) + lane * DPL;      // 16 positions of a page are contiguous
        const __half* vb1 = last >= 16 ? kv.at_slot(slot, layer, 1, kvh, p0 + 16) + lane * DPL : vb0;
        constexpr int NBV = (DPL == 4 && G <= 4) ? 16 : 8;                          // V rows in flight per lane (registers permitting)
#pragma unroll
        for (int pp0 = 0; pp0 < 32; pp0 += NBV) {
            if (pp0 > last) break;
            uint2 vv[NBV];
#pragma unroll
            for (int j = 0; j < NBV; j++) {
                const int pp = min(pp0 + j, last);
                const __half* vp = (pp < 16 ? vb0 : vb1) + (size_t)(pp & 15) * HD;
                if (DPL == 4) vv[j] = *(const uint2*)vp; else { vv[j].x = *(const uint32_t*)vp; vv[j].y = 0; }
            }
#pragma unroll
            for (int j = 0; j < NBV; j++) {
                const float2 a = h2_to_f2(vv[j].x), b = h2_to_f2(vv[j].y);
                const float v[4] = {a.x, a.y, b.x, b.y};
#pragma unroll
                for (int g = 0; g < G; g++) {
                    const float pr = __shfl_sync(0xffffffffu, pe[g], pp0 + j);
#pragma unroll
                    for (int i = 0; i < DPL; i++) acc[g][i] += pr * v[i];
                }
            }
        }
    }
#pragma unroll
    for (int g = 0; g < G; g++) {
        if (lane == 0) { sm_m[warp][g] = mx[g]; sm_l[warp][g] = l[g]; }
#pragma unroll
        for (int i = 0; i < DPL; i++) sm_acc[warp][g][lane * DPL + i] = acc[g][i];
    }
    __syncthreads();
    for (int idx = tid; idx < G * HD; idx += NW * 32) {
        const int g = idx / HD, d = idx % HD;
        float mm = sm_m[0][g];
#pragma unroll
        for (int w = 1; w < NW; w++) mm = fmaxf(mm, sm_m[w][g]);
        float num = 0.f, den = 0.f;
#pragma unroll
        for (int w = 0; w < NW; w++) { const float e = __expf(sm_m[w][g] - mm); num += sm_acc[w][g][d] * e; den += sm_l[w][g] * e; }
        ((unsigned short*)O)[(size_t)m * n_head * HD + (kvh * G + g) * HD + (perm ? kperm4(d) : d)] = pack16(num / den, perm);
    }
}

// next[m] = argmax_v logits[m][v] (lowest index wins ties); one block per row.  The vocabulary may be split into n_parts
// slices of V columns (tensor-parallel lm_head): slice p of row m is at logits + p * part_stride + m * ld, its first
// column is token p * V.
__global__ void argmax_rows_kernel(const float* __restrict__ logits, int V, int ld, int n_parts, size_t part_stride, int* __restrict__ next) {
    __shared__ float sv[32]; __shared__ int si[32];
    pdl_launch_dependents(); pdl_wait();
    float bv = -INFINITY; int bi = 0x7fffffff;
    const int V4 = V & ~3;                                           // rows start 16-byte aligned (ld % 4 == 0)
    for (int p = 0; p < n_parts; p++) {
        const float* row = logits + (size_t)p * part_stride + (size_t)blockIdx.x * ld; const int base = p * V;
#pragma unroll 4
        for (int i = threadIdx.x * 4; i < V4; i += blockDim.x * 4) {
            const float4 v = *(const float4*)(row + i);
            if (v.x > bv) { bv = v.x; bi = base + i; } if (v.y > bv) { bv = v.y; bi = base + i + 1; } if (v.z > bv) { bv = v.z; bi = base + i + 2; } if (v.w > bv) { bv = v.w; bi = base + i + 3; }
        }
        for (int i = V4 + threadIdx.x; i < V; i += blockDim.x) { const float v = row[i]; if (v > bv || (v == bv && base + i < bi)) { bv = v; bi = base + i; } }
    }
    for (int o = 16; o; o >>= 1) { const float v = __shfl_xor_sync(0xffffffffu, bv, o); const int ix = __shfl_xor_sync(0xffffffffu, bi, o); if (v > bv || (v == bv && ix < bi)) { bv = v; bi = ix; } }
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < (int)(blockDim.x >> 5); w++) if (sv[w] > bv || (sv[w] == bv && si[w] < bi)) { bv = sv[w]; bi = si[w]; }
        next[blockIdx.x] = bi < n_parts * V ? bi : 0;                  // all-NaN logits: stay inside the vocabulary
    }
}

// ACT[m][i] = bf16(silu(GU[m][2i]) * GU[m][2i+1])   (gate/up rows are interleaved in the fused matrix)
__global__ void swiglu_rows_kernel(const float* __restrict__ GU, int M, int F, __nv_bfloat16* __restrict__ ACT, bool perm, const __grid_constant__ QgSplit sp) {
    pdl_launch_dependents(); pdl_wait();
    // four outputs (eight interleaved gate/up columns: one tile, one partial lookup) per thread; F % 4 == 0
    const long long n4 = (long long)M * (F / 4);
    for (long long i4 = blockIdx.x * (long long)blockDim.x + threadIdx.x; i4 < n4; i4 += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(i4 / (F / 4)), f0 = (int)(i4 % (F / 4)) * 4;
        const float4* src = (const float4*)(GU + (size_t)m * 2 * F + 2 * f0);
        const float4 a = src[0], b = src[1];
        float t[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        if (sp.ws) qg_gather<8>(sp, m, 2 * f0, t);
        float o[4];
#pragma unroll
        for (int k = 0; k < 4; k++) o[k] = t[2 * k] / (1.f + __expf(-t[2 * k])) * t[2 * k + 1];
        uint2 w = perm ? make_uint2(pack16x2(o[0], o[2], true), pack16x2(o[1], o[3], true)) : make_uint2(pack16x2(o[0], o[1], false), pack16x2(o[2], o[3], false));
        *(uint2*)(ACT + (size_t)m * F + f0) = w;
    }
}

}  // namespace b200
