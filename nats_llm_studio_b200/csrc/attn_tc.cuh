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
// attn_decode_tc_kernel: the same two MMAs for ONE new position per sequence.  One CTA per (token m, kv head); the query
// tile is the G heads that share the kv head — rows m * n_head + kvh * G .. of Q viewed as [token * head][head_dim]
// (tmap_qd is that view, box of 8 rows) — all at the token's position.  Built to run TWO CTAs per SM, since a CTA walks
// its KV tiles serially and the step is bound by how many bytes an SM keeps in flight:
//   * the query tile is the 8-row swizzle atom per 64-column panel (1 KB); the MMA's other 120 A rows read whatever
//     follows in shared memory (finite garbage), their accumulator rows are never read;
//   * K and V are single-buffered (32 KB each at head size 128): K(j+1) is requested as soon as S(j) has been computed,
//     V(j+1) as soon as O += P.V (j) has — each load runs under the other operand's work;
//   * P overwrites S in tensor memory (pass 2 of the softmax writes columns [c/2, c/2+16) after reading [c, c+32)), so
//     S|P + O fit 256 columns and two CTAs share the SM's 512;
//   * only warp 0 holds real rows (G <= 8 of its lanes): it alone runs the softmax; a second warp issues the TMA boxes
//     (one per lane) and the MMAs; the sequence's block table is read into shared memory once, in one round trip.
// Position and slot come from the step's device-side row map (the step is a CUDA graph: nothing position-dependent
// may sit in the launch parameters).  Output: fp16 in kperm4 order, the activation format of qgemm_kernel.
template <int HD> struct FdCfg {
    static constexpr int PANELS = HD / 64;
    static constexpr uint32_t Q_BYTES = PANELS * 1024, KV_BYTES = FA_BN * HD * 2;          // 8 query rows | one K or V tile
    // the A descriptor of the last panel reads 16 KB from its start: K's buffer follows Q, so that stays inside the allocation
    static constexpr size_t SMEM = 1024 + 1024 * ((Q_BYTES + 1023) / 1024) + 2 * (size_t)KV_BYTES + 64 + BTAB_MAX * 4 + 64;   // + barriers + block table
    static constexpr uint32_t TM_S = 0, TM_O = 128, TM_COLS = 256;
};

constexpr int FD_THREADS = 64;      // warp 0: softmax of the G real rows (TMEM lanes 0..31); warp 1: TMA loads and MMA issue
template <int HD>
__global__ void __launch_bounds__(FD_THREADS, 2)
attn_decode_tc_kernel(const __grid_constant__ CUtensorMap tmap_qd, const __grid_constant__ CUtensorMap tmap_kv, RowMap rm, int G, int layer,
                      int n_head, int n_head_kv, float kq_scale, KvLayout kv, __half* __restrict__ O) {
    using C = FdCfg<HD>;
    extern __shared__ uint8_t fd_smem_raw[];
    uint8_t* base = (uint8_t*)(((uintptr_t)fd_smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = base; uint8_t* sK = base + 1024 * ((C::Q_BYTES + 1023) / 1024); uint8_t* sV = sK + C::KV_BYTES;
    uint64_t* bars = (uint64_t*)(sV + C::KV_BYTES);
    uint64_t *q_full = bars, *k_full = bars + 1, *v_full = bars + 2, *s_done = bars + 3, *o_done = bars + 4, *p_full = bars + 5; uint32_t* tmem_slot = (uint32_t*)(bars + 6);
    int* s_bt = (int*)(bars + 8);                                      // this sequence's block table (pages of 16 positions)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m = blockIdx.x, kvh = blockIdx.y;
    if (tid == 0) { for (int i = 0; i < 6; i++) mbar_init(&bars[i], 1); fence_barrier_init(); }
    if (warp == 0) tcgen05_alloc(tmem_slot, C::TM_COLS);
    pdl_launch_dependents(); pdl_wait();                              // Q and the new K/V rows come from the preceding kernels
    const int row_t = m * n_head + kvh * G, pos = rm.p(m), slot = rm.s(m);
    const int n_kv_tiles = pos / FA_BN + 1;
    for (int i = tid; i <= (pos >> 4); i += FD_THREADS) s_bt[i] = kv.block_table[(size_t)slot * kv.pages_per_seq + i];   // one round trip, not one per page
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem = *tmem_slot;
    const float sl2 = kq_scale * 1.4426950408889634f;                 // scores are kept in log2 units: exp2f((s - m) * sl2)

    if (warp == 1) {
        // ===== loads and MMAs.  A tile = 8 pages x PANELS boxes of 64 columns x 16 positions; lane i issues box i =====
        auto load_tile = [&](int j, int which, uint8_t* dst, uint64_t* bar) {      // which: 0 = K, 1 = V
            if (lane == 0) mbar_expect_tx(bar, C::KV_BYTES);
            __syncwarp();
            for (int b = lane; b < (FA_BN / 16) * C::PANELS; b += 32) {
                const int pg = b / C::PANELS, pn = b % C::PANELS;
                const int p = min(j * FA_BN + pg * 16, pos);       // pages beyond the position: re-read a valid page (masked in the softmax)
                const int r = (int)(((((size_t)layer * kv.n_pages + s_bt[p >> 4]) * 2 + which) * kv.n_head_kv + kvh) * 16);
                tma_load_2d(dst + (size_t)pn * (FA_BN * 128) + (size_t)pg * 2048, &tmap_kv, pn * 64, r, bar);
            }
        };
        if (lane == 0) { mbar_expect_tx(q_full, C::Q_BYTES); for (int pn = 0; pn < C::PANELS; pn++) tma_load_2d(sQ + (size_t)pn * 1024, &tmap_qd, pn * 64, row_t, q_full); }
        load_tile(0, 0, sK, k_full); load_tile(0, 1, sV, v_full);
        for (int j = 0; j < n_kv_tiles; j++) {
            const uint32_t ph = j & 1;
            if (lane == 0) {
                if (j == 0) mbar_wait(q_full, 0);
                mbar_wait(k_full, ph);
                tcgen05_fence_after();
                constexpr uint32_t idesc = fa_idesc(FA_BM, FA_BN, false);
#pragma unroll
                for (int k = 0; k < HD / 16; k++) {                    // S = Q . K^T: 16 head-dim columns per instruction
                    const uint64_t da = umma_desc_sw128(sQ + (size_t)(k / 4) * 1024) + 2 * (k % 4);
                    const uint64_t db = umma_desc_sw128(sK + (size_t)(k / 4) * (FA_BN * 128)) + 2 * (k % 4);
                    tcgen05_mma_f16(tmem + C::TM_S, da, db, idesc, k ? 1u : 0u);
                }
                tcgen05_commit(s_done);
                mbar_wait(s_done, ph);                                 // S(j) has read K: the buffer takes the next tile
            }
            __syncwarp();
            if (j + 1 < n_kv_tiles) load_tile(j + 1, 0, sK, k_full);
            if (lane == 0) {
                mbar_wait(p_full, ph);                                 // the softmax warp has written P(j) (and rescaled O)
                mbar_wait(v_full, ph);
                tcgen05_fence_after();
                constexpr uint32_t idesc = fa_idesc(FA_BM, HD, true);
#pragma unroll
                for (int k = 0; k < FA_BN / 16; k++) {                 // O += P . V: 16 positions per instruction
                    const uint64_t db = umma_desc_mn_sw128(sV + (size_t)k * 2048, FA_BN * 128);
                    tcgen05_mma_f16_ts(tmem + C::TM_O, tmem + C::TM_S + k * 8, db, idesc, (j | k) ? 1u : 0u);
                }
                tcgen05_commit(o_done);
                mbar_wait(o_done, ph);                                 // P.V(j) has read V (and P: S(j+1) may overwrite it)
            }
            __syncwarp();
            if (j + 1 < n_kv_tiles) load_tile(j + 1, 1, sV, v_full);
        }
    } else {
        // ===== softmax: lane r owns query row r (rows >= G are other heads' queries or zero fill: computed, never stored) =====
        float m_run = -INFINITY, l_run = 0.f;
        for (int j = 0; j < n_kv_tiles; j++) {
            mbar_wait(s_done, j & 1);                                  // commit order: S(j) done implies P.V(j-1) done, so O may be rescaled
            __syncwarp();
            tcgen05_fence_after();
            const int kv0 = j * FA_BN;
            float m_new = m_run;
#pragma unroll 1
            for (int c = 0; c < FA_BN; c += 32) {
                uint32_t sc[32]; tmem_ld_x32(tmem + C::TM_S + c, sc);
#pragma unroll
                for (int i = 0; i < 32; i++) if (kv0 + c + i <= pos) m_new = fmaxf(m_new, __uint_as_float(sc[i]));      // positions past the token's own are masked
            }
            const float corr = m_run == -INFINITY ? 0.f : exp2f((m_run - m_new) * sl2);
            float l_add = 0.f;
#pragma unroll 1
            for (int c = 0; c < FA_BN; c += 32) {                      // P overwrites S: columns [c/2, c/2+16) <- scores [c, c+32), all of them read before
                uint32_t sc[32], pk[16]; tmem_ld_x32(tmem + C::TM_S + c, sc);
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    const float p0 = kv0 + c + i <= pos ? exp2f((__uint_as_float(sc[i]) - m_new) * sl2) : 0.f;
                    const float p1 = kv0 + c + i + 1 <= pos ? exp2f((__uint_as_float(sc[i + 1]) - m_new) * sl2) : 0.f;
                    const __half2 h = __floats2half2_rn(p0, p1);
                    const float2 back = __half22float2(h); l_add += back.x + back.y;      // the sum uses the rounded values the MMA will see
                    pk[i / 2] = *(const uint32_t*)&h;
                }
                tmem_st_x8(tmem + C::TM_S + c / 2, pk); tmem_st_x8(tmem + C::TM_S + c / 2 + 8, pk + 8);
            }
            l_run = l_run * corr + l_add; m_run = m_new;
            if (j > 0) {                                               // O *= corr, in place
#pragma unroll 1
                for (int c = 0; c < HD; c += 32) {
                    uint32_t o[32]; tmem_ld_x32(tmem + C::TM_O + c, o);
#pragma unroll
                    for (int i = 0; i < 32; i++) o[i] = __float_as_uint(__uint_as_float(o[i]) * corr);
                    tmem_st_x32(tmem + C::TM_O + c, o);
                }
            }
            tmem_st_wait();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
        }
        // ---- epilogue: rows of Q in the [token * head][head_dim] view are rows of O in the same view
        mbar_wait(o_done, (n_kv_tiles - 1) & 1);
        __syncwarp();
        tcgen05_fence_after();
        const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
#pragma unroll 1
        for (int c = 0; c < HD; c += 32) {
            uint32_t o[32]; tmem_ld_x32(tmem + C::TM_O + c, o);
            if (lane < G) {
                __half* dst = O + (size_t)(row_t + lane) * HD + c;
#pragma unroll
                for (int i = 0; i < 32; i += 8) {                      // kperm4: (0, 2, 1, 3) inside every aligned group of four
                    uint4 w;
                    const __half2 t0 = __floats2half2_rn(clamp_h(__uint_as_float(o[i]) * inv), clamp_h(__uint_as_float(o[i + 2]) * inv)), t1 = __floats2half2_rn(clamp_h(__uint_as_float(o[i + 1]) * inv), clamp_h(__uint_as_float(o[i + 3]) * inv));
                    const __half2 t2 = __floats2half2_rn(clamp_h(__uint_as_float(o[i + 4]) * inv), clamp_h(__uint_as_float(o[i + 6]) * inv)), t3 = __floats2half2_rn(clamp_h(__uint_as_float(o[i + 5]) * inv), clamp_h(__uint_as_float(o[i + 7]) * inv));
                    w.x = *(const uint32_t*)&t0; w.y = *(const uint32_t*)&t1; w.z = *(const uint32_t*)&t2; w.w = *(const uint32_t*)&t3;
                    *(uint4*)(dst + i) = w;
                }
            }
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 0) tcgen05_dealloc(tmem, C::TM_COLS);
}

}  // namespace b200
