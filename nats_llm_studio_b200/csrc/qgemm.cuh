// qgemm.cuh — dequant-in-GEMM for batched decode (BASELINE.json "bs64"):
//     C[b][n] (+)= scale * sum_k X[b][k] * W[n][k],   b < B <= 64,  W quantised (Q4_K / Q5_K / Q6_K / Q8_0)
// The weights are read ONCE per step in their quantised form (the op is HBM-bound at B <= 64),
// expanded to fp16 in shared memory by 16 dequant warps, and contracted on the tensor cores:
// tcgen05.mma kind::f16, A = 128 weight rows (fp16, 128-byte-swizzled K-major tiles written by the dequant
// warps), B = the 64 activation rows (fp16, TMA), D = fp32 in TMEM (two 64-column accumulators, ping-pong).
//
// Tile layout ("TL", built once from the streaming layout by tile_repack_kernel): the matrix is cut into tiles of
// 128 rows and K-steps of KSTEP elements (256 for the K-quants = one super-block, 128 for Q8_0); the bytes of one
// (tile, K-step) are one contiguous blob = one bulk copy, ordered inside so that 32 lanes = 32 consecutive rows
// read their operands with conflict-free 128-bit shared loads:
//   Q4_K (18432 B): hdr[row][16] (d, dmin, 12 engine-form scale bytes) | qs[j2 4][h 2][row][16]
//   Q5_K (22528 B): hdr[row][16] | qh[h 2][row][16] | qs[j2 4][h 2][row][16]
//   Q6_K (26880 B): ql[n 2][piece 4][row][16] | qh[n 2][h 2][row][16] | sc[n,h 4][row][4] | d[row][2]
//   Q8_0 (17408 B): q[kb 2][piece 4][row][16] | d[kb 2][row][2 x fp16]
// Work split: stream-K.  The (tile, K-step) units of all matrices of one launch form one sequence that is cut into
// n_sm equal contiguous ranges, one per CTA; a tile whose K range ends up in several CTAs is summed, in CTA order
// (deterministic), by qgemm_fixup_kernel from fp32 partials.
// K order: inside every aligned group of 4 K positions the A tile holds elements in the order (0, 2, 1, 3) — that is
// how nibble pairs fall out of a 32-bit word without a permute — so X must be stored with the same permutation
// (kperm4); the row kernels that produce X for this path do that.
// Numerics: weights are expanded as fp16(16 * w) (two or three fp16 roundings, rel. 2^-11 each), activations fp16
// (clamped to the finite range by the row kernels that write them), fp32 accumulation; the factor 16 keeps small scales out of the fp16 subnormal range and is removed in the epilogue.
#pragma once
#include "prefill.cuh"

namespace b200 {

constexpr int QG_ROWS = 128, QG_NB = 64;
constexpr int QG_DQ_WARPS = 16, QG_FIRST_DQ = 3, QG_FIRST_EPI = QG_FIRST_DQ + QG_DQ_WARPS, QG_WARPS = QG_FIRST_EPI + 4, QG_THREADS = QG_WARPS * 32;
constexpr int QG_A_SLOTS = 4, QG_X_SLOTS = 8, QG_MAX_RAW = 8;
constexpr uint32_t QG_A_BYTES = QG_ROWS * 128, QG_X_BYTES = QG_NB * 128;
constexpr float QG_PRESCALE = 16.f;

__host__ __device__ constexpr int qg_kstep(int t) { return t == T_Q8_0 ? 128 : 256; }
__host__ __device__ constexpr int qg_blob_bytes(int t) {
    return t == T_Q4_K ? 128 * 144 : t == T_Q5_K ? 128 * 176 : t == T_Q6_K ? 128 * 210 : t == T_Q8_0 ? 128 * 136 : 0;
}
__host__ __device__ inline size_t tile_bytes(int t, int64_t N, int K) { return (size_t)((N + QG_ROWS - 1) / QG_ROWS) * (K / qg_kstep(t)) * qg_blob_bytes(t); }

// blob byte o -> (row in tile, raw block inside the K-step, raw byte inside that block)
__host__ __device__ inline void tile_src(int t, int o, int& row, int& blk, int& i) {
    blk = 0;
    if (t == T_Q4_K || t == T_Q5_K) {
        if (o < 2048) { row = o / 16; i = o % 16; return; }
        o -= 2048;
        if (t == T_Q5_K) { if (o < 4096) { row = (o % 2048) / 16; i = 16 + 16 * (o / 2048) + o % 16; return; } o -= 4096; }
        row = (o % 2048) / 16; i = (t == T_Q4_K ? 16 : 48) + 16 * (o / 2048) + o % 16; return;
    }
    if (t == T_Q6_K) {
        if (o < 16384) { row = (o % 2048) / 16; i = 16 * (o / 2048) + o % 16; return; }                 // ql[64n + 16pc + within]
        o -= 16384;
        if (o < 8192) { row = (o % 2048) / 16; i = 128 + 16 * (o / 2048) + o % 16; return; }            // qh[32n + 16h + within]
        o -= 8192;
        if (o < 2048) { const int p = o / 512, n = p / 2, h = p % 2; row = (o % 512) / 4; i = 192 + 8 * n + 2 * (o % 4) + h; return; }
        o -= 2048; row = o / 2; i = 208 + o % 2; return;
    }
    // Q8_0, K-step of 128 = 4 blocks
    if (o < 16384) { const int pc = o / 2048, e = 16 * pc + o % 16; row = (o % 2048) / 16; blk = e / 32; i = 2 + e % 32; return; }
    o -= 16384; { const int kb = o / 512, b = o % 4; row = (o % 512) / 4; blk = 2 * kb + b / 2; i = b % 2; }
}

// streaming layout (quant.cuh) -> tile layout; rows >= N read as zero bytes
__global__ void tile_repack_kernel(const uint8_t* __restrict__ w, int type, int64_t N, int K, uint8_t* __restrict__ out) {
    const int ks = K / qg_kstep(type), blob = qg_blob_bytes(type), ng = n_groups(K), gb = group_bytes(type);
    const int64_t n_tiles = (N + QG_ROWS - 1) / QG_ROWS, total = n_tiles * ks * (int64_t)blob;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int o = (int)(idx % blob); const int64_t u = idx / blob; const int kstep = (int)(u % ks); const int64_t tile = u / ks;
        int row, blk, i; tile_src(type, o, row, blk, i);
        const int64_t r = tile * QG_ROWS + row;
        uint8_t v = 0;
        if (r < N) {
            const int sb = type == T_Q8_0 ? kstep * 4 + blk : kstep;          // raw block index inside the row
            const int per_grp = type == T_Q8_0 ? 64 : 8, g = sb / per_grp;
            v = w[(((r >> 1) * ng + g) * 2 + (r & 1)) * (size_t)gb + dst_byte(type, sb % per_grp, i)];
        }
        out[idx] = v;
    }
}

struct QSeg { const uint8_t* tl; float* C; int type, n_tiles, n_rows, ks, step0, tile0; };
struct QGemmParams {
    QSeg seg[3]; int n_seg, total_steps, n_tiles, B, ldc, mode, n_raw; float scale; float* ws; uint32_t slot_bytes;
    int debug;      // profiling only (B200_QG_DEBUG): 1 = skip the expansion math, 2 = skip the MMAs
};

__device__ __forceinline__ int qg_seg_of_step(const QGemmParams& P, int s) { int i = 0; while (i + 1 < P.n_seg && s >= P.seg[i + 1].step0) i++; return i; }

// ------------------------------------------------------------------------------------ fp16x2 helpers
__device__ __forceinline__ uint32_t h2fma(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t h2sub(uint32_t a, uint32_t b) { uint32_t d; asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t h2mul(uint32_t a, uint32_t b) { uint32_t d; asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t h2dup(float v) { uint32_t d; asm("cvt.rn.f16x2.f32 %0, %1, %1;" : "=r"(d) : "f"(v)); return d; }
__device__ __forceinline__ float h16f(uint32_t bits) { return __half2float(__ushort_as_half((unsigned short)bits)); }
__device__ __forceinline__ uint4 lds128(const uint8_t* p) { return *(const uint4*)p; }
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
// 16-byte chunk c (8 K positions) of row `row` inside a 128-row x 64-K swizzled tile
__device__ __forceinline__ uint32_t a_chunk(uint32_t tile, int row, int c) { return tile + (row >> 3) * 1024 + (row & 7) * 128 + ((c ^ (row & 7)) << 4); }
__device__ __forceinline__ void mbar_arrive_n(uint64_t* bar, uint32_t n) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(n) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

constexpr uint32_t H2_MAGIC = 0x64006400u /*1024*/, H2_M1024 = 0xE400E400u, H2_1152 = 0x64806480u, H2_1056 = 0x64206420u;

// ------------------------------------------------------------------------------------ per-type expansion, one lane = one row
// Q4_K / Q5_K: lane expands the 64 elements of chunk j2 (sub-blocks 2*j2, 2*j2+1) of its row into k-block tile `tile`
template <bool Q5>
__device__ __forceinline__ void dq_k45(const uint8_t* raw, int row, int j2, uint32_t tile) {
    const uint4 hd = lds128(raw + row * 16);
    const uint32_t f = j2 == 0 ? hd.y : j2 == 1 ? __funnelshift_r(hd.y, hd.z, 24) : j2 == 2 ? __funnelshift_r(hd.z, hd.w, 16) : hd.w >> 8;
    const float d = h16f(hd.x & 0xFFFF) * QG_PRESCALE, dmin = h16f(hd.x >> 16) * QG_PRESCALE;
    const uint32_t S0 = h2dup(d * (float)(f & 63)), S1 = h2dup(d * (float)((f >> 6) & 63) * 0.0625f);
    const uint32_t M0 = h2dup(dmin * (float)((f >> 12) & 63)), M1 = h2dup(dmin * (float)((f >> 18) & 63));
    const uint32_t C0 = h2mul(S0, H2_M1024), C1 = h2mul(S1, H2_M1024);
    const uint8_t* qs = raw + 2048 + (Q5 ? 4096 : 0) + (size_t)(j2 * 2) * 2048 + row * 16;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint4 q = lds128(qs + h * 2048);
        uint4 hq = make_uint4(0, 0, 0, 0);
        if (Q5) hq = lds128(raw + 2048 + h * 2048 + row * 16);
        const uint32_t xs[4] = {q.x, q.y, q.z, q.w}, hs[4] = {hq.x, hq.y, hq.z, hq.w};
        uint32_t lo[8], hi[8];
#pragma unroll
        for (int w = 0; w < 4; w++) {
            const uint32_t x = xs[w], y = x >> 8;
            uint32_t l02 = (x & 0x000F000Fu) | H2_MAGIC, h02 = (x & 0x00F000F0u) | H2_MAGIC, l13 = (y & 0x000F000Fu) | H2_MAGIC, h13 = (y & 0x00F000F0u) | H2_MAGIC;
            if (Q5) {
                const uint32_t a = hs[w] >> (2 * j2), b = a >> 8;
                l02 |= (a << 4) & 0x00100010u; h02 |= (a << 7) & 0x01000100u; l13 |= (b << 4) & 0x00100010u; h13 |= (b << 7) & 0x01000100u;
            }
            lo[2 * w] = h2sub(h2fma(l02, S0, C0), M0); lo[2 * w + 1] = h2sub(h2fma(l13, S0, C0), M0);
            hi[2 * w] = h2sub(h2fma(h02, S1, C1), M1); hi[2 * w + 1] = h2sub(h2fma(h13, S1, C1), M1);
        }
        sts128(a_chunk(tile, row, 2 * h), lo[0], lo[1], lo[2], lo[3]); sts128(a_chunk(tile, row, 2 * h + 1), lo[4], lo[5], lo[6], lo[7]);
        sts128(a_chunk(tile, row, 4 + 2 * h), hi[0], hi[1], hi[2], hi[3]); sts128(a_chunk(tile, row, 5 + 2 * h), hi[4], hi[5], hi[6], hi[7]);
    }
}

// Q8_0: lane expands block (2*kb + half) = 32 elements of its row into chunks 4*half .. 4*half+3 of the k-block tile
__device__ __forceinline__ void dq_q80(const uint8_t* raw, int row, int kb, int half, uint32_t tile) {
    const uint32_t D = h2dup(h16f(*(const unsigned short*)(raw + 16384 + ((kb * 128 + row) * 2 + half) * 2)) * QG_PRESCALE);
#pragma unroll
    for (int pc = 0; pc < 2; pc++) {
        const uint4 q = lds128(raw + (size_t)((kb * 4 + 2 * half + pc) * 128 + row) * 16);
        const uint32_t xs[4] = {q.x, q.y, q.z, q.w}; uint32_t o[8];
#pragma unroll
        for (int w = 0; w < 4; w++) {
            const uint32_t u = xs[w] ^ 0x80808080u;
            o[2 * w] = h2mul(h2sub(__byte_perm(u, 0x64646464u, 0x4240), H2_1152), D);
            o[2 * w + 1] = h2mul(h2sub(__byte_perm(u, 0x64646464u, 0x4341), H2_1152), D);
        }
        const int c = 4 * half + 2 * pc;
        sts128(a_chunk(tile, row, c), o[0], o[1], o[2], o[3]); sts128(a_chunk(tile, row, c + 1), o[4], o[5], o[6], o[7]);
    }
}

// Q6_K: lane expands, for 128-half n and 16-run h, the 64 elements (grp 0..3) x (l in [16h, 16h+16)) of its row:
// grp 0,1 -> k-block 2n (tile0), grp 2,3 -> k-block 2n+1 (tile1)
__device__ __forceinline__ void dq_q6k(const uint8_t* raw, int row, int n, int h, uint32_t tile0, uint32_t tile1) {
    const uint4 qa = lds128(raw + (size_t)((n * 4 + h) * 128 + row) * 16), qb = lds128(raw + (size_t)((n * 4 + 2 + h) * 128 + row) * 16);
    const uint4 qh = lds128(raw + 16384 + (size_t)((n * 2 + h) * 128 + row) * 16);
    const uint32_t sc = *(const uint32_t*)(raw + 24576 + ((n * 2 + h) * 128 + row) * 4);
    const float d = h16f(*(const unsigned short*)(raw + 26624 + row * 2)) * QG_PRESCALE;
    const uint32_t S0 = h2dup(d * (float)(int8_t)(sc & 0xFF)), S1 = h2dup(d * (float)(int8_t)((sc >> 8) & 0xFF)),
                   S2 = h2dup(d * (float)(int8_t)((sc >> 16) & 0xFF)), S3 = h2dup(d * (float)(int8_t)(sc >> 24));
    const uint32_t as[4] = {qa.x, qa.y, qa.z, qa.w}, bs[4] = {qb.x, qb.y, qb.z, qb.w}, hs[4] = {qh.x, qh.y, qh.z, qh.w};
    uint32_t g0[8], g1[8], g2[8], g3[8];
#pragma unroll
    for (int w = 0; w < 4; w++) {
        const uint32_t xa = as[w], xb = bs[w], hq = hs[w];
        const uint32_t m4 = 0x000F000Fu, m2 = 0x00300030u;
        g0[2 * w]     = h2mul(h2sub(((xa)       & m4) | ((hq << 4) & m2) | H2_MAGIC, H2_1056), S0);
        g0[2 * w + 1] = h2mul(h2sub(((xa >> 8)  & m4) | ((hq >> 4) & m2) | H2_MAGIC, H2_1056), S0);
        g1[2 * w]     = h2mul(h2sub(((xb)       & m4) | ((hq << 2) & m2) | H2_MAGIC, H2_1056), S1);
        g1[2 * w + 1] = h2mul(h2sub(((xb >> 8)  & m4) | ((hq >> 6) & m2) | H2_MAGIC, H2_1056), S1);
        g2[2 * w]     = h2mul(h2sub(((xa >> 4)  & m4) | ((hq)      & m2) | H2_MAGIC, H2_1056), S2);
        g2[2 * w + 1] = h2mul(h2sub(((xa >> 12) & m4) | ((hq >> 8) & m2) | H2_MAGIC, H2_1056), S2);
        g3[2 * w]     = h2mul(h2sub(((xb >> 4)  & m4) | ((hq >> 2) & m2) | H2_MAGIC, H2_1056), S3);
        g3[2 * w + 1] = h2mul(h2sub(((xb >> 12) & m4) | ((hq >> 10) & m2) | H2_MAGIC, H2_1056), S3);
    }
    sts128(a_chunk(tile0, row, 2 * h), g0[0], g0[1], g0[2], g0[3]);     sts128(a_chunk(tile0, row, 2 * h + 1), g0[4], g0[5], g0[6], g0[7]);
    sts128(a_chunk(tile0, row, 4 + 2 * h), g1[0], g1[1], g1[2], g1[3]); sts128(a_chunk(tile0, row, 5 + 2 * h), g1[4], g1[5], g1[6], g1[7]);
    sts128(a_chunk(tile1, row, 2 * h), g2[0], g2[1], g2[2], g2[3]);     sts128(a_chunk(tile1, row, 2 * h + 1), g2[4], g2[5], g2[6], g2[7]);
    sts128(a_chunk(tile1, row, 4 + 2 * h), g3[0], g3[1], g3[2], g3[3]); sts128(a_chunk(tile1, row, 5 + 2 * h), g3[4], g3[5], g3[6], g3[7]);
}

__host__ __device__ constexpr uint32_t umma_idesc_f16(int M, int N) {
    return (1u << 4) /*D: F32*/ | (0u << 7) /*A: F16*/ | (0u << 10) /*B: F16*/ | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld_32x32b_x32_nowait(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
          "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
          "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}

// shared-memory carve-up (all offsets from the 1024-aligned base)
struct QgSmem {
    uint8_t *a, *x, *raw; uint64_t *raw_full, *raw_empty, *x_full, *x_empty, *a_full, *a_empty, *acc_full, *acc_empty; uint32_t* tmem_slot;
};
constexpr uint32_t QG_BAR_BYTES = 512;
__host__ __device__ inline size_t qg_smem_bytes(uint32_t slot_bytes, int n_raw) {
    return 1024 + QG_A_SLOTS * QG_A_BYTES + QG_X_SLOTS * QG_X_BYTES + (size_t)n_raw * slot_bytes + QG_BAR_BYTES;
}

// ------------------------------------------------------------------------------------ the kernel
__global__ void __launch_bounds__(QG_THREADS, 1)
qgemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ QGemmParams P) {
    extern __shared__ uint8_t qg_smem_raw[];
    QgSmem S;
    {
        uint8_t* base = (uint8_t*)(((uintptr_t)qg_smem_raw + 1023) & ~(uintptr_t)1023);
        S.a = base; S.x = S.a + QG_A_SLOTS * QG_A_BYTES; S.raw = S.x + QG_X_SLOTS * QG_X_BYTES;
        uint64_t* b = (uint64_t*)(S.raw + (size_t)P.n_raw * P.slot_bytes);
        S.raw_full = b; S.raw_empty = b + QG_MAX_RAW; S.x_full = b + 2 * QG_MAX_RAW; S.x_empty = S.x_full + QG_X_SLOTS;
        S.a_full = S.x_empty + QG_X_SLOTS; S.a_empty = S.a_full + QG_A_SLOTS; S.acc_full = S.a_empty + QG_A_SLOTS; S.acc_empty = S.acc_full + 2;
        S.tmem_slot = (uint32_t*)(S.acc_empty + 2);
    }
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int i = 0; i < P.n_raw; i++) { mbar_init(&S.raw_full[i], 1); mbar_init(&S.raw_empty[i], QG_DQ_WARPS); }
        for (int i = 0; i < QG_X_SLOTS; i++) { mbar_init(&S.x_full[i], 1); mbar_init(&S.x_empty[i], 1); }
        for (int i = 0; i < QG_A_SLOTS; i++) { mbar_init(&S.a_full[i], QG_DQ_WARPS); mbar_init(&S.a_empty[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&S.acc_full[i], 1); mbar_init(&S.acc_empty[i], 4); }
        fence_barrier_init();
    }
    if (warp == 2) tcgen05_alloc(S.tmem_slot, 2 * QG_NB);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem = *S.tmem_slot;

    const int G = gridDim.x, cta = blockIdx.x;
    const int s_begin = (int)((long long)cta * P.total_steps / G), s_end = (int)((long long)(cta + 1) * P.total_steps / G);

    if (warp == 0) {                                   // ===== weight producer: one bulk copy per (tile, K-step) blob =====
        if (lane == 0) {
            int it = 0;
            for (int s = s_begin; s < s_end; s++, it++) {
                const QSeg& sg = P.seg[qg_seg_of_step(P, s)];
                const int rs = it % P.n_raw; const uint32_t ph = (it / P.n_raw) & 1, bytes = qg_blob_bytes(sg.type);
                mbar_wait(&S.raw_empty[rs], ph ^ 1);
                if (P.debug & 8) { mbar_arrive(&S.raw_full[rs]); continue; }
                mbar_expect_tx(&S.raw_full[rs], bytes);
                bulk_g2s(S.raw + (size_t)rs * P.slot_bytes, sg.tl + (size_t)(s - sg.step0) * bytes, bytes, &S.raw_full[rs]);
            }
        }
    } else if (warp == 1) {                            // ===== activation producer: one 64 x 64 bf16 box per k-block =====
        if (lane == 0) {
            int kbc = 0;
            for (int s = s_begin; s < s_end; s++) {
                const QSeg& sg = P.seg[qg_seg_of_step(P, s)];
                const int kstep = (s - sg.step0) % sg.ks, nkb = qg_kstep(sg.type) / 64;
                for (int kb = 0; kb < nkb; kb++, kbc++) {
                    const int xs = kbc % QG_X_SLOTS; const uint32_t ph = (kbc / QG_X_SLOTS) & 1;
                    mbar_wait(&S.x_empty[xs], ph ^ 1);
                    if (P.debug & 4) { mbar_arrive(&S.x_full[xs]); continue; }
                    mbar_expect_tx(&S.x_full[xs], QG_X_BYTES);
                    tma_load_2d(S.x + xs * QG_X_BYTES, &tmap_x, (kstep * nkb + kb) * 64, 0, &S.x_full[xs]);
                }
            }
        }
    } else if (warp == 2) {                            // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(QG_ROWS, QG_NB);
            int kbc = 0, piece = 0, s = s_begin;
            while (s < s_end) {
                const QSeg& sg = P.seg[qg_seg_of_step(P, s)];
                const int k0 = (s - sg.step0) % sg.ks, n_st = min(sg.ks - k0, s_end - s), nkb = qg_kstep(sg.type) / 64;
                const int buf = piece & 1;
                mbar_wait(&S.acc_empty[buf], ((piece >> 1) & 1) ^ 1);
                tcgen05_fence_after();
                for (int i = 0; i < n_st * nkb; i++, kbc++) {
                    const int as = kbc % QG_A_SLOTS, xs = kbc % QG_X_SLOTS;
                    mbar_wait(&S.a_full[as], (kbc / QG_A_SLOTS) & 1);
                    mbar_wait(&S.x_full[xs], (kbc / QG_X_SLOTS) & 1);
                    tcgen05_fence_after();
                    const uint64_t da = umma_desc_sw128(S.a + as * QG_A_BYTES), db = umma_desc_sw128(S.x + xs * QG_X_BYTES);
#pragma unroll
                    for (int k = 0; k < 4; k++) if (!(P.debug & 2)) tcgen05_mma_f16(tmem + buf * QG_NB, da + 2 * k, db + 2 * k, idesc, (i | k) ? 1u : 0u);
                    tcgen05_commit(&S.a_empty[as]);
                    tcgen05_commit(&S.x_empty[xs]);
                }
                tcgen05_commit(&S.acc_full[buf]);
                s += n_st; piece++;
            }
        }
    } else if (warp < QG_FIRST_EPI) {                  // ===== dequant warps: raw blob -> fp16 swizzled k-block tiles =====
        const int w = warp - QG_FIRST_DQ, row = (w & 3) * 32 + lane, p = w >> 2;
        int it = 0, kbc = 0;
        for (int s = s_begin; s < s_end; s++, it++) {
            const int type = P.seg[qg_seg_of_step(P, s)].type;
            const int rs = it % P.n_raw;
            const uint8_t* raw = S.raw + (size_t)rs * P.slot_bytes;
            mbar_wait(&S.raw_full[rs], (it / P.n_raw) & 1);
            if (type == T_Q8_0) {
                const int kb = p >> 1, as = (kbc + kb) % QG_A_SLOTS;
                mbar_wait(&S.a_empty[as], (((kbc + kb) / QG_A_SLOTS) & 1) ^ 1);
                if (!(P.debug & 1)) dq_q80(raw, row, kb, p & 1, smem_u32(S.a + as * QG_A_BYTES));
                fence_proxy_async_smem(); __syncwarp();
                if (lane == 0) mbar_arrive_n(&S.a_full[as], 2);
                kbc += 2;
            } else if (type == T_Q6_K) {
                const int n = p >> 1, a0 = (kbc + 2 * n) % QG_A_SLOTS, a1 = (kbc + 2 * n + 1) % QG_A_SLOTS;
                mbar_wait(&S.a_empty[a0], (((kbc + 2 * n) / QG_A_SLOTS) & 1) ^ 1);
                mbar_wait(&S.a_empty[a1], (((kbc + 2 * n + 1) / QG_A_SLOTS) & 1) ^ 1);
                if (!(P.debug & 1)) dq_q6k(raw, row, n, p & 1, smem_u32(S.a + a0 * QG_A_BYTES), smem_u32(S.a + a1 * QG_A_BYTES));
                fence_proxy_async_smem(); __syncwarp();
                if (lane == 0) { mbar_arrive_n(&S.a_full[a0], 2); mbar_arrive_n(&S.a_full[a1], 2); }
                kbc += 4;
            } else {
                const int as = (kbc + p) % QG_A_SLOTS;
                mbar_wait(&S.a_empty[as], (((kbc + p) / QG_A_SLOTS) & 1) ^ 1);
                if (P.debug & 1) {}
                else if (type == T_Q4_K) dq_k45<false>(raw, row, p, smem_u32(S.a + as * QG_A_BYTES));
                else dq_k45<true>(raw, row, p, smem_u32(S.a + as * QG_A_BYTES));
                fence_proxy_async_smem(); __syncwarp();
                if (lane == 0) mbar_arrive_n(&S.a_full[as], 4);
                kbc += 4;
            }
            if (lane == 0) mbar_arrive(&S.raw_empty[rs]);       // after __syncwarp: every lane has finished reading the blob
        }
    } else {                                           // ===== epilogue warps: TMEM -> C (whole tile) or fp32 partial =====
        const int q = warp & 3;                        // TMEM lane quadrant this warp may read
        int piece = 0, s = s_begin;
        const float scale = P.scale * (1.f / QG_PRESCALE);
        while (s < s_end) {
            const QSeg& sg = P.seg[qg_seg_of_step(P, s)];
            const int tl = (s - sg.step0) / sg.ks, k0 = (s - sg.step0) % sg.ks, n_st = min(sg.ks - k0, s_end - s);
            const int buf = piece & 1;
            mbar_wait(&S.acc_full[buf], (piece >> 1) & 1);
            tcgen05_fence_after();
            uint32_t r[64];
            const uint32_t ta = tmem + buf * QG_NB + ((uint32_t)(q * 32) << 16);
            tmem_ld_32x32b_x32_nowait(ta, r); tmem_ld_32x32b_x32_nowait(ta + 32, r + 32);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&S.acc_empty[buf]);      // registers hold the tile: the accumulator may be reused
            const int n = tl * QG_ROWS + q * 32 + lane;
            if (n_st == sg.ks) {
                if (n < sg.n_rows) {
                    float* dst = sg.C + n;
#pragma unroll
                    for (int b = 0; b < QG_NB; b++)
                        if (b < P.B) { float v = __uint_as_float(r[b]) * scale; if (P.mode == 1) v += dst[(size_t)b * P.ldc]; dst[(size_t)b * P.ldc] = v; }
                }
            } else {
                const int first_tile = qg_seg_of_step(P, s_begin) == qg_seg_of_step(P, s) && (s_begin - sg.step0) / sg.ks == tl;
                float* dst = P.ws + ((size_t)(cta * 2 + (first_tile ? 0 : 1)) * QG_NB) * QG_ROWS + q * 32 + lane;
#pragma unroll
                for (int b = 0; b < QG_NB; b++) if (b < P.B) dst[(size_t)b * QG_ROWS] = __uint_as_float(r[b]);
            }
            s += n_st; piece++;
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 2) tcgen05_dealloc(tmem, 2 * QG_NB);
}

// Sum the fp32 partials of every tile whose K range was cut across CTAs (ascending CTA = ascending K: deterministic)
__global__ void qgemm_fixup_kernel(const __grid_constant__ QGemmParams P, int G) {
    const int t = blockIdx.x;                          // global tile index
    int si = 0; while (si + 1 < P.n_seg && t >= P.seg[si + 1].tile0) si++;
    const QSeg& sg = P.seg[si];
    const int tl = t - sg.tile0, s0 = sg.step0 + tl * sg.ks, s1 = s0 + sg.ks - 1;
    const long long T = P.total_steps;
    const int c_lo = (int)((((long long)s0 + 1) * G + T - 1) / T) - 1, c_hi = (int)((((long long)s1 + 1) * G + T - 1) / T) - 1;
    if (c_lo == c_hi) return;                          // whole tile in one CTA: written directly
    const int n = tl * QG_ROWS + threadIdx.x;
    if (n >= sg.n_rows) return;
    const float scale = P.scale * (1.f / QG_PRESCALE);
    constexpr int BB = 8;                              // batch rows per block (blockIdx.y): independent loads in flight
    const int b0 = blockIdx.y * BB;
    float acc[BB];
#pragma unroll
    for (int i = 0; i < BB; i++) acc[i] = 0.f;
    for (int c = c_lo; c <= c_hi; c++) {
        const int cb = (int)((long long)c * T / G), ce = (int)((long long)(c + 1) * T / G);
        if (cb == ce) continue;                                             // CTA without work (fewer steps than CTAs)
        const int slot = cb >= s0 ? 0 : 1;                                  // slot 0: the tile is the first one CTA c touches
        const float* src = P.ws + ((size_t)(c * 2 + slot) * QG_NB + b0) * QG_ROWS + threadIdx.x;
#pragma unroll
        for (int i = 0; i < BB; i++) if (b0 + i < P.B) acc[i] += src[(size_t)i * QG_ROWS];
    }
#pragma unroll
    for (int i = 0; i < BB; i++) {
        if (b0 + i >= P.B) break;
        float v = acc[i] * scale; float* dst = sg.C + (size_t)(b0 + i) * P.ldc + n;
        if (P.mode == 1) v += *dst;
        *dst = v;
    }
}

}  // namespace b200
