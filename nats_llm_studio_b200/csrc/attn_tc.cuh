// attn_tc.cuh — K7: causal prompt attention on the 5th-generation tensor cores (flash-style), for the rows of a prompt
// chunk (runs of consecutive positions of one sequence each, cut into query tiles) over the paged fp16 KV cache.
// Part of what LM Studio's /api/v0/chat/completions does behind /root/reference/nats_llm_studio.go:158-179.
//
// One CTA per (query tile of <= 128 rows of one run, head), 128 threads.  Per 128-position KV tile:
//   S = Q . K^T        tcgen05.mma kind::f16, Q and K tiles staged by TMA (K-major, 128-byte swizzle), S in tensor memory
//   softmax            thread r owns query row r (= TMEM lane r): tcgen05.ld of its 128 scores, running max / sum, the
//                      probabilities go BACK to tensor memory as fp16 (tcgen05.st) — P never touches shared memory
//   O += P . V         tcgen05.mma with A = P from tensor memory and B = the V tile exactly as it lies in the cache
//                      ([position][head_dim], i.e. MN-major for this product: no transpose pass); O in tensor memory,
//                      rescaled in place when the running max moves.
// K and V tiles are double-buffered: the TMA loads of tile j+1 are in flight during tile j's two MMAs and softmax.
// KV pages are 16 positions: a tile is 8 (page, 16-row) boxes per 64-column panel, addressed through the block table.
#pragma once
#include "prefill.cuh"
#include "qgemm.cuh"

namespace b200 {

constexpr int FA_BM = 128, FA_BN = 128, FA_THREADS = 128;
template <int HD> struct FaCfg {
    static constexpr int PANELS = HD / 64;                          // 64-column (128-byte) swizzle panels per row
    static constexpr uint32_t TILE_BYTES = FA_BN * HD * 2;          // one Q / K / V tile
    static constexpr size_t SMEM = 5 * (size_t)TILE_BYTES + 1024 + 128;       // Q + 2 x (K, V)
    static constexpr uint32_t TM_S = 0, TM_O = 128, TM_P = 128 + HD, TM_COLS = 512;
};
// instruction descriptor, fp16 x fp16 -> fp32; b_mn: B operand is MN-major (V: head_dim contiguous, positions are K)
__host__ __device__ constexpr uint32_t fa_idesc(int M, int N, bool b_mn) {
    return (1u << 4) | (0u << 7) | (0u << 10) | ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// MN-major operand tile [64-element panels along MN][rows along K of 128 bytes], 128-byte swizzle: 8-row groups 1024 B apart
// (stride byte offset), panels `panel_bytes` apart (leading byte offset)
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(const void* smem, uint32_t panel_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_u32(smem) & 0x3FFFF) >> 4);
    d |= (uint64_t)((panel_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ void tmem_ld_x32(uint32_t taddr, uint32_t* r) { tmem_ld_32x32b_x32_nowait(taddr, r); asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// One launch covers every run of a prompt chunk: blockIdx.x indexes a table of query tiles {first row in the chunk, rows in
// the tile (<= 128), position of the first row, sequence slot}.  (One launch per run made the choice of kernel — and with
// it the rounding of a request's logits — depend on what else was packed into the chunk.)
constexpr int FA_MAX_TILES = 64;
struct FaTiles { int4 t[FA_MAX_TILES]; };

template <int HD>
__global__ void __launch_bounds__(FA_THREADS, 1)
attn_prefill_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv, const __grid_constant__ FaTiles tiles, int layer,
                       int n_head, int n_head_kv, float kq_scale, KvLayout kv, __nv_bfloat16* __restrict__ O) {
    using C = FaCfg<HD>;
    extern __shared__ uint8_t fa_smem_raw[];
    uint8_t* base = (uint8_t*)(((uintptr_t)fa_smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = base; uint8_t* sK = base + C::TILE_BYTES; uint8_t* sV = sK + 2 * C::TILE_BYTES;
    uint64_t* bars = (uint64_t*)(base + 5 * (size_t)C::TILE_BYTES);
    uint64_t *q_full = bars, *k_full = bars + 1 /*[2]*/, *v_full = bars + 3 /*[2]*/, *s_done = bars + 5, *o_done = bars + 6; uint32_t* tmem_slot = (uint32_t*)(bars + 8);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int head = blockIdx.y, kvh = head / (n_head / n_head_kv);
    if (tid == 0) { for (int i = 0; i < 7; i++) mbar_init(&bars[i], 1); fence_barrier_init(); }
    if (warp == 0) tcgen05_alloc(tmem_slot, C::TM_COLS);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem = *tmem_slot;
    pdl_launch_dependents(); pdl_wait();                              // Q and the new K/V rows come from the preceding kernels
    const int4 tile = tiles.t[blockIdx.x];
    const int row_t = tile.x, rows_here = tile.y, pos_t = tile.z, slot = tile.w;
    const int last_pos = pos_t + rows_here - 1;                       // newest position any row of this tile may see
    const int n_kv_tiles = last_pos / FA_BN + 1;

    // one KV tile: 8 pages x PANELS boxes of (64 columns x 16 positions) for K, the same for V
    auto load_kv = [&](int j, int buf) {
        mbar_expect_tx(&k_full[buf], C::TILE_BYTES); mbar_expect_tx(&v_full[buf], C::TILE_BYTES);
        for (int pg = 0; pg < FA_BN / 16; pg++) {
            const int p = min(j * FA_BN + pg * 16, last_pos);       // pages beyond the last position: re-read a valid page (masked below)
            const int page = kv.block_table[(size_t)slot * kv.pages_per_seq + (p >> 4)];
            const int rk = (int)(((((size_t)layer * kv.n_pages + page) * 2 + 0) * kv.n_head_kv + kvh) * 16), rv = rk + kv.n_head_kv * 16;
            for (int pn = 0; pn < C::PANELS; pn++) {
                tma_load_2d(sK + (size_t)buf * C::TILE_BYTES + (size_t)pn * (FA_BN * 128) + (size_t)pg * 2048, &tmap_kv, pn * 64, rk, &k_full[buf]);
                tma_load_2d(sV + (size_t)buf * C::TILE_BYTES + (size_t)pn * (FA_BN * 128) + (size_t)pg * 2048, &tmap_kv, pn * 64, rv, &v_full[buf]);
            }
        }
    };
    if (tid == 0) {
        mbar_expect_tx(q_full, C::TILE_BYTES);
        for (int pn = 0; pn < C::PANELS; pn++) tma_load_2d(sQ + (size_t)pn * (FA_BM * 128), &tmap_q, head * HD + pn * 64, row_t, q_full);
        load_kv(0, 0);
    }
    const int my_pos = pos_t + tid;                                   // position of this thread's query row
    const float sl2 = kq_scale * 1.4426950408889634f;                 // scores are kept in log2 units: exp2f((s - m) * sl2)
    float m_run = -INFINITY, l_run = 0.f;
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;            // this warp's quarter of the TMEM lanes
    for (int j = 0; j < n_kv_tiles; j++) {
        const int buf = j & 1; const uint32_t ph = (j >> 1) & 1;
        if (tid == 0) {
            if (j + 1 < n_kv_tiles) {                                  // the other buffer's previous tenant is tile j-1: its P.V must have read V
                if (j > 0) mbar_wait(o_done, (j - 1) & 1);
                load_kv(j + 1, buf ^ 1);
            }
            if (j == 0) mbar_wait(q_full, 0);
            mbar_wait(&k_full[buf], ph);
            tcgen05_fence_after();
            constexpr uint32_t idesc = fa_idesc(FA_BM, FA_BN, false);
#pragma unroll
            for (int k = 0; k < HD / 16; k++) {                        // S = Q . K^T: 16 head-dim columns per instruction
                const uint64_t da = umma_desc_sw128(sQ + (size_t)(k / 4) * (FA_BM * 128)) + 2 * (k % 4);
                const uint64_t db = umma_desc_sw128(sK + (size_t)buf * C::TILE_BYTES + (size_t)(k / 4) * (FA_BN * 128)) + 2 * (k % 4);
                tcgen05_mma_f16(tmem + C::TM_S, da, db, idesc, k ? 1u : 0u);
            }
            tcgen05_commit(s_done);
        }
        // ---- softmax of this thread's row
        mbar_wait(s_done, j & 1);
        __syncwarp();
        tcgen05_fence_after();
        {
        // two passes over the row's 128 scores, 32 at a time (registers): the maximum, then the probabilities
        const int kv0 = j * FA_BN;
        float m_new = m_run;
#pragma unroll 1
        for (int c = 0; c < FA_BN; c += 32) {
            uint32_t sc[32]; tmem_ld_x32(tmem + C::TM_S + lane_off + c, sc);
#pragma unroll
            for (int i = 0; i < 32; i++) if (kv0 + c + i <= my_pos) m_new = fmaxf(m_new, __uint_as_float(sc[i]));      // causal mask (also covers positions past the end)
        }
        // rows beyond n_rows (tile padding) see garbage queries: harmless, never stored.  Every real row sees position 0.
        const float corr = m_run == -INFINITY ? 0.f : exp2f((m_run - m_new) * sl2);
        // the previous tile's P.V must have finished before P is overwritten and O rescaled
        if (j > 0) { mbar_wait(o_done, (j - 1) & 1); tcgen05_fence_after(); }
        float l_add = 0.f;
#pragma unroll 1
        for (int c = 0; c < FA_BN; c += 32) {
            uint32_t sc[32], pk[16]; tmem_ld_x32(tmem + C::TM_S + lane_off + c, sc);
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
                const float p0 = kv0 + c + i <= my_pos ? exp2f((__uint_as_float(sc[i]) - m_new) * sl2) : 0.f;
                const float p1 = kv0 + c + i + 1 <= my_pos ? exp2f((__uint_as_float(sc[i + 1]) - m_new) * sl2) : 0.f;
                const __half2 h = __floats2half2_rn(p0, p1);
                const float2 back = __half22float2(h); l_add += back.x + back.y;      // the sum uses the rounded values the MMA will see
                pk[i / 2] = *(const uint32_t*)&h;
            }
            tmem_st_x8(tmem + C::TM_P + lane_off + c / 2, pk); tmem_st_x8(tmem + C::TM_P + lane_off + c / 2 + 8, pk + 8);
        }
        l_run = l_run * corr + l_add; m_run = m_new;
        if (j > 0) {                                                   // O *= corr, in place
#pragma unroll 1
            for (int c = 0; c < HD; c += 32) {
                uint32_t o[32]; tmem_ld_x32(tmem + C::TM_O + lane_off + c, o);
#pragma unroll
                for (int i = 0; i < 32; i++) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
                tmem_st_x32(tmem + C::TM_O + lane_off + c, o);
            }
        }
        }
        tmem_st_wait();
        tcgen05_fence_before();
        __syncthreads();
        if (tid == 0) {
            tcgen05_fence_after();
            mbar_wait(&v_full[buf], ph);
            tcgen05_fence_after();
            constexpr uint32_t idesc = fa_idesc(FA_BM, HD, true);
#pragma unroll
            for (int k = 0; k < FA_BN / 16; k++) {                     // O += P . V: 16 positions per instruction
                const uint64_t db = umma_desc_mn_sw128(sV + (size_t)buf * C::TILE_BYTES + (size_t)k * 2048, FA_BN * 128);
                tcgen05_mma_f16_ts(tmem + C::TM_O, tmem + C::TM_P + k * 8, db, idesc, (j | k) ? 1u : 0u);
            }
            tcgen05_commit(o_done);
        }
    }
    // ---- epilogue: O / l -> bf16 row of the output [row][head * HD + d]
    mbar_wait(o_done, (n_kv_tiles - 1) & 1);
    __syncwarp();
    tcgen05_fence_after();
    const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
    {
    const int row = row_t + tid;
#pragma unroll 1
    for (int c = 0; c < HD; c += 32) {
        uint32_t o[32]; tmem_ld_x32(tmem + C::TM_O + lane_off + c, o);
        if (tid < rows_here) {
            __nv_bfloat16* dst = O + (size_t)row * n_head * HD + head * HD + c;
#pragma unroll
            for (int i = 0; i < 32; i += 8) {
                uint4 w;
                __nv_bfloat162 t0 = __floats2bfloat162_rn(__uint_as_float(o[i]) * inv, __uint_as_float(o[i + 1]) * inv), t1 = __floats2bfloat162_rn(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
                __nv_bfloat162 t2 = __floats2bfloat162_rn(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv), t3 = __floats2bfloat162_rn(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
                w.x = *(uint32_t*)&t0; w.y = *(uint32_t*)&t1; w.z = *(uint32_t*)&t2; w.w = *(uint32_t*)&t3;
                *(uint4*)(dst + i) = w;
            }
        }
    }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) tcgen05_dealloc(tmem, C::TM_COLS);
}

// ------------------------------------------------------------------------------------ batched decode step
// attn_decode_tc_kernel: attention for ONE new position per sequence, one CTA per (token m, kv head), with the MMA operands
// SWAPPED so that every tensor-memory lane holds real data:
//   S^T = K . Q^T     A = the K tile (M = 128 rows, FD_BN of them positions of the tile), B = the G query heads of the kv group padded to N = 16 rows
//                     (rows m * n_head + kvh * G .. of Q viewed as [token * head][head_dim]); lane t = position t
//   softmax           thread t owns position t: G scores from one tcgen05.ld, the per-head maximum by warp shuffles and one
//                     block barrier, G exponentials, P^T written to shared memory (fp16, 128-byte swizzle, K-major B operand)
//   O^T += V^T . P    A = the V tile as it lies in the cache ([position][head_dim] = M-major A), B = P^T; lane d = output
//                     dimension d, 16 columns = heads; rescaled in place per column when a head's maximum moves
// 32 TMEM columns and 50 KB of shared memory per CTA (80-position tiles, below): four CTAs per SM cover each other's serial
// load -> S -> softmax -> P.V chain.  (The query-rows-as-M form, G real rows of 128 and one warp doing the whole softmax, measured 5.4 - 6.5 ms per
// B = 64 step against 4.81 for the CUDA-core kernel.)  Position and slot come from the step's device-side row map (the
// step is a CUDA graph).  Output: fp16 in kperm4 order, the activation format of qgemm_kernel.  Head size 128 only.
constexpr int FD_THREADS = 128, FD_N = 16, FD_BTAB = 256;          // FD_BTAB pages = 4096 positions per sequence
// Positions per KV tile.  80 (not 128): 40 KB of K + V per CTA lets FOUR CTAs share an SM, and 64 x 8 = 512 (token, kv head)
// items then run in one wave of 592 slots; with 128-position tiles (three CTAs per SM, 444 slots) the 68 items of the second
// round ran alone at a fraction of the bandwidth: 40 % of the kernel's time.  The S^T MMA still has M = 128: its rows 80..127
// read on into the next panel / the V buffer (finite garbage) and are masked.  One tile size for every batch size, so that a
// row's rounding does not depend on its batch.
#ifndef B200_FD_BN
#define B200_FD_BN 80
#endif
constexpr int FD_BN = B200_FD_BN;
static_assert(FD_BN % 16 == 0 && FD_BN >= 64 && FD_BN <= 128, "decode attention tile: whole pages, at least one 64-position panel");
struct FdCfg {
    static constexpr uint32_t KV_BYTES = FD_BN * 128 * 2, QP_BYTES = 2 * FD_N * 128;        // one K or V tile | Q or P^T: 2 panels x 16 rows x 128 B
    static constexpr size_t SMEM = 1024 + 2 * (size_t)QP_BYTES + 2 * (size_t)KV_BYTES + 64 + 4 * FD_N * 4 + FD_BTAB * 4;
    static constexpr uint32_t TM_S = 0, TM_O = FD_N, TM_COLS = 32;
};
__host__ __device__ constexpr uint32_t fd_idesc(int M, int N, bool a_mn) {      // fp16 x fp16 -> fp32; a_mn: A operand is M-major
    return (1u << 4) | (0u << 7) | (0u << 10) | ((a_mn ? 1u : 0u) << 15) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t* r) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n\ttcgen05.wait::ld.sync.aligned;"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__global__ void __launch_bounds__(FD_THREADS, FD_BN <= 80 ? 4 : 3)
attn_decode_tc_kernel(const __grid_constant__ CUtensorMap tmap_qd, const __grid_constant__ CUtensorMap tmap_kv, RowMap rm, int G, int layer,
                      int n_head, int n_head_kv, float kq_scale, KvLayout kv, __half* __restrict__ O) {
    using C = FdCfg; constexpr int HD = 128;
    extern __shared__ uint8_t fd_smem_raw[];
    uint8_t* base = (uint8_t*)(((uintptr_t)fd_smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = base; uint8_t* sP = sQ + C::QP_BYTES; uint8_t* sK = sP + C::QP_BYTES; uint8_t* sV = sK + C::KV_BYTES;
    uint64_t* bars = (uint64_t*)(sV + C::KV_BYTES);
    uint64_t *q_full = bars, *k_full = bars + 1, *v_full = bars + 2, *s_done = bars + 3, *o_done = bars + 4; uint32_t* tmem_slot = (uint32_t*)(bars + 6);
    float* red = (float*)(bars + 8);                                   // [4 warps][FD_N]
    int* s_bt = (int*)(red + 4 * FD_N);                                // this sequence's block table (pages of 16 positions)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m = blockIdx.x, kvh = blockIdx.y;
    if (tid == 0) { for (int i = 0; i < 5; i++) mbar_init(&bars[i], 1); fence_barrier_init(); }
    if (warp == 0) tcgen05_alloc(tmem_slot, C::TM_COLS);
    for (int i = tid; i < (int)C::QP_BYTES / 16; i += FD_THREADS) ((uint4*)sP)[i] = make_uint4(0, 0, 0, 0);      // rows >= G of P^T stay zero
    pdl_launch_dependents(); pdl_wait();                              // Q and the new K/V rows come from the preceding kernels
    const int row_t = m * n_head + kvh * G, pos = rm.p(m), slot = rm.s(m);
    const int n_kv_tiles = pos / FD_BN + 1;
    for (int i = tid; i <= (pos >> 4); i += FD_THREADS) s_bt[i] = kv.block_table[(size_t)slot * kv.pages_per_seq + i];   // one round trip, not one per page
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem = *tmem_slot;
    const uint32_t lane_off = (uint32_t)(warp * 32) << 16;            // this warp's quarter of the TMEM lanes
    const float sl2 = kq_scale * 1.4426950408889634f;                 // scores are kept in log2 units: exp2f((s - m) * sl2)

    // a tile = 8 pages x 2 boxes of 64 columns x 16 positions; lane i of warp 0 issues box i
    auto load_tile = [&](int j, int which, uint8_t* dst, uint64_t* bar) {      // which: 0 = K, 1 = V; called by all of warp 0
        if (lane == 0) mbar_expect_tx(bar, C::KV_BYTES);
        __syncwarp();
        if (lane < 2 * (FD_BN / 16)) {
            const int pg = lane >> 1, pn = lane & 1;
            const int p = min(j * FD_BN + pg * 16, pos);           // pages beyond the position: re-read a valid page (masked in the softmax)
            const int r = (int)(((((size_t)layer * kv.n_pages + s_bt[p >> 4]) * 2 + which) * kv.n_head_kv + kvh) * 16);
            tma_load_2d(dst + (size_t)pn * (FD_BN * 128) + (size_t)pg * 2048, &tmap_kv, pn * 64, r, bar);
        }
    };
    if (warp == 0) {
        if (lane == 0) { mbar_expect_tx(q_full, C::QP_BYTES); for (int pn = 0; pn < 2; pn++) tma_load_2d(sQ + (size_t)pn * (FD_N * 128), &tmap_qd, pn * 64, row_t, q_full); }
        load_tile(0, 0, sK, k_full); load_tile(0, 1, sV, v_full);
    }
    float m_run[8], l_part[8];
#pragma unroll
    for (int g = 0; g < 8; g++) { m_run[g] = -INFINITY; l_part[g] = 0.f; }
    for (int j = 0; j < n_kv_tiles; j++) {
        const uint32_t ph = j & 1;
        if (warp == 0) {
            if (lane == 0) {
                if (j == 0) mbar_wait(q_full, 0);
                mbar_wait(k_full, ph);
                tcgen05_fence_after();
                constexpr uint32_t idesc = fd_idesc(128, FD_N, false);       // M = 128 whatever the tile: rows >= FD_BN are masked
#pragma unroll
                for (int k = 0; k < HD / 16; k++) {                    // S^T = K . Q^T: 16 head-dim columns per instruction
                    const uint64_t da = umma_desc_sw128(sK + (size_t)(k / 4) * (FD_BN * 128)) + 2 * (k % 4);
                    const uint64_t db = umma_desc_sw128(sQ + (size_t)(k / 4) * (FD_N * 128)) + 2 * (k % 4);
                    tcgen05_mma_f16(tmem + C::TM_S, da, db, idesc, k ? 1u : 0u);
                }
                tcgen05_commit(s_done);
                if (j > 0) mbar_wait(o_done, (j - 1) & 1);             // P.V(j-1) has read V: the buffer takes this tile's V
            }
            __syncwarp();
            if (j > 0) load_tile(j, 1, sV, v_full);
        }
        mbar_wait(s_done, ph);
        __syncwarp();
        tcgen05_fence_after();
        if (warp == 0 && j + 1 < n_kv_tiles) load_tile(j + 1, 0, sK, k_full);      // S^T(j) has read K
        // ---- softmax: this thread's position against the G heads
        uint32_t sc[16]; tmem_ld_x16(tmem + C::TM_S + lane_off, sc);
        const bool ok = tid < FD_BN && j * FD_BN + tid <= pos;         // rows past the tile and positions past the token's own are masked
        float mx[8];
#pragma unroll
        for (int g = 0; g < 8; g++) {
            float v = (ok && g < G) ? __uint_as_float(sc[g]) : -INFINITY;
#pragma unroll
            for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
            mx[g] = v;
        }
#pragma unroll
        for (int g = 0; g < 8; g++) if (lane == g) red[warp * FD_N + g] = mx[g];        // (mx[lane] would put the array in local memory)
        if (j > 0) { mbar_wait(o_done, (j - 1) & 1); tcgen05_fence_after(); }       // P.V(j-1) has read P^T and written O^T
        __syncthreads();
        float corr[8];
        const int panel = tid >> 6; const uint32_t colb = (uint32_t)(tid & 63) * 2;
#pragma unroll
        for (int g = 0; g < 8; g++) if (g < G) {
            const float mt = fmaxf(fmaxf(red[g], red[FD_N + g]), fmaxf(red[2 * FD_N + g], red[3 * FD_N + g]));
            const float mn = fmaxf(m_run[g], mt);                      // finite: every tile holds at least position kv0 <= pos
            corr[g] = m_run[g] == -INFINITY ? 0.f : exp2f((m_run[g] - mn) * sl2);
            const __half pe = __float2half_rn(ok ? exp2f((__uint_as_float(sc[g]) - mn) * sl2) : 0.f);
            l_part[g] = l_part[g] * corr[g] + __half2float(pe);        // the sum uses the rounded value the MMA will see
            m_run[g] = mn;
            *(__half*)(sP + panel * (FD_N * 128) + g * 128 + (((colb >> 4) ^ (uint32_t)(g & 7)) << 4) + (colb & 15)) = pe;
        } else corr[g] = 1.f;
        if (j > 0) {                                                   // O^T[d][g] *= corr[g], in place; this thread's d = tid
            uint32_t o[16]; tmem_ld_x16(tmem + C::TM_O + lane_off, o);
#pragma unroll
            for (int g = 0; g < 8; g++) o[g] = __float_as_uint(__uint_as_float(o[g]) * corr[g]);
            tmem_st_x8(tmem + C::TM_O + lane_off, o);
            tmem_st_wait();
        }
        fence_proxy_async_smem();                                      // P^T was written with ordinary stores; the MMA reads it through the async proxy
        tcgen05_fence_before();
        __syncthreads();
        if (tid == 0) {
            tcgen05_fence_after();
            mbar_wait(v_full, ph);
            tcgen05_fence_after();
            constexpr uint32_t idesc = fd_idesc(HD, FD_N, true);
#pragma unroll
            for (int k = 0; k < FD_BN / 16; k++) {                     // O^T += V^T . P: 16 positions per instruction
                const uint64_t da = umma_desc_mn_sw128(sV + (size_t)k * 2048, FD_BN * 128);
                const uint64_t db = umma_desc_sw128(sP + (size_t)(k / 4) * (FD_N * 128)) + 2 * (k % 4);
                tcgen05_mma_f16(tmem + C::TM_O, da, db, idesc, (j | k) ? 1u : 0u);
            }
            tcgen05_commit(o_done);
        }
    }
    // ---- epilogue: total l per head (sum over the 128 positions' partial sums), then O^T[d][g] / l[g] -> O[(row_t + g)][d]
#pragma unroll
    for (int g = 0; g < 8; g++) {
        float v = l_part[g];
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        l_part[g] = v;
    }
    __syncthreads();                                                   // red[] is free (every thread has read the last tile's maxima)
#pragma unroll
    for (int g = 0; g < 8; g++) if (lane == g) red[warp * FD_N + g] = l_part[g];
    mbar_wait(o_done, (n_kv_tiles - 1) & 1);
    tcgen05_fence_after();
    __syncthreads();
    uint32_t o[16]; tmem_ld_x16(tmem + C::TM_O + lane_off, o);
    const int d = tid, dp = kperm4(d);
#pragma unroll
    for (int g = 0; g < 8; g++) if (g < G) {
        const float l = red[g] + red[FD_N + g] + red[2 * FD_N + g] + red[3 * FD_N + g];
        O[(size_t)(row_t + g) * HD + dp] = __float2half_rn(clamp_h(l > 0.f ? __uint_as_float(o[g]) / l : 0.f));
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) tcgen05_dealloc(tmem, C::TM_COLS);
}

}  // namespace b200
