// qgemm.cuh — dequant-in-GEMM for batched decode (BASELINE.json "bs64"):
//     C[b][n] (+)= scale * sum_k X[b][k] * W[n][k],   b < B <= 64,  W quantised (Q4_K / Q5_K / Q6_K / Q8_0)
// The weights are read ONCE per step in their quantised form (the op is HBM-bound at B <= 64),
// expanded to fp16 by 16 dequant warps and contracted on the tensor cores: tcgen05.mma kind::f16 with
//   A = 128 weight rows, fp16, written by the dequant warps straight into TENSOR MEMORY (tcgen05.st; lane = row,
//       one 32-bit column = two K positions) — a ring of 8 k-blocks (2 K-steps) in TMEM columns 128..383,
//   B = the 64 activation rows (fp16, TMA, 128-byte swizzle, shared memory),
//   D = fp32 accumulators in TMEM columns 0..127 (two of 64 columns, ping-pong between tiles).
// Keeping A out of shared memory leaves ~160 KB for the raw-weight ring (bulk copies in flight) and lets the
// expansion run two K-steps ahead of the MMAs.
//
// Tile layout ("TL", built once from the streaming layout by tile_repack_kernel): the matrix is cut into tiles of
// 128 rows and K-steps of 256 elements; the bytes of one (tile, K-step) are one contiguous blob = one bulk copy,
// ordered inside so that 32 lanes = 32 consecutive rows read their operands with conflict-free 128-bit shared loads:
//   Q4_K (18432 B): hdr[row][16] (d, dmin, 12 engine-form scale bytes) | qs[j2 4][h 2][row][16]
//   Q5_K (22528 B): hdr[row][16] | qh[h 2][row][16] | qs[j2 4][h 2][row][16]
//   Q6_K (26880 B): ql[n 2][piece 4][row][16] | qh[n 2][h 2][row][16] | sc[n,h 4][row][4] | d[row][2]
//   Q8_0 (34816 B): q[kb 4][piece 4][row][16] | d[kb 4][row][2 x fp16]
// Work split: stream-K.  The (tile, K-step) units of all matrices of one launch form one sequence that is cut into
// equal contiguous ranges, one per CTA; a tile whose K range ends up in several CTAs is left as fp32 partials that the
// consuming row kernel (RoPE, SwiGLU, the next RMSNorm) sums in CTA order on the fly — deterministic, no atomics, no
// reduction pass (QgSplit in prefill.cuh).  Matrices with many more tiles than CTAs (lm_head) use tile-aligned ranges.
// K order: inside every aligned group of 4 K positions the A operand holds elements in the order (0, 2, 1, 3) — that
// is how nibble pairs fall out of a 32-bit word without a permute — so X must be stored with the same permutation
// (kperm4); the row kernels that produce X for this path do that.
// Numerics: weights are expanded as fp16(16 * w) (two or three fp16 roundings, rel. 2^-11 each), activations fp16
// (clamped to the finite range by the row kernels that write them), fp32 accumulation; the factor 16 keeps small
// scales out of the fp16 subnormal range and is removed in the epilogue.
#pragma once
#include "prefill.cuh"

namespace b200 {

constexpr int QG_ROWS = 128, QG_NB = 64, QG_KSTEP = 256;
constexpr int QG_DQ_WARPS = 16, QG_FIRST_DQ = 3, QG_FIRST_EPI = QG_FIRST_DQ + QG_DQ_WARPS, QG_WARPS = QG_FIRST_EPI + 4, QG_THREADS = QG_WARPS * 32;
// X and A rings have the same depth: one tcgen05.commit per K-step releases both (a second commit per step cost ~0.2 us).
// Every ring depth is EVEN: the two expansion groups take alternate K-steps, so with an even depth a ring slot (and its
// barriers) is always used by the same group, which then sees every phase of them; with an odd depth a group comes back to
// a barrier only every second phase, and a parity wait cannot tell phase n from phase n+2 (seen as sporadic launch failures
// with a 3-deep Q8_0 weight ring).
constexpr int QG_A_STEPS = 2, QG_X_STEPS = QG_A_STEPS, QG_MAX_RAW = 8, QG_MAX_TILES = 4096;
constexpr uint32_t QG_X_BYTES = 4 * QG_NB * 128;              // one K-step of activations: 4 k-blocks of 64 rows x 128 B
constexpr uint32_t QG_TMEM_COLS = 512, QG_TMEM_A0 = 2 * QG_NB; // accumulators in columns [0,128), A ring in [128,512)
constexpr float QG_PRESCALE = 16.f;

__host__ __device__ constexpr int qg_blob_bytes(int t) {
    return t == T_Q4_K ? 128 * 144 : t == T_Q5_K ? 128 * 176 : t == T_Q6_K ? 128 * 210 : t == T_Q8_0 ? 128 * 272 : 0;
}
__host__ __device__ inline size_t tile_bytes(int t, int64_t N, int K) { return (size_t)((N + QG_ROWS - 1) / QG_ROWS) * (K / QG_KSTEP) * qg_blob_bytes(t); }

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
    // Q8_0: 8 blocks of 32 per K-step
    if (o < 32768) { const int e = 16 * (o / 2048) + o % 16; row = (o % 2048) / 16; blk = e / 32; i = 2 + e % 32; return; }
    o -= 32768; { const int kb = o / 512, b = o % 4; row = (o % 512) / 4; blk = 2 * kb + b / 2; i = b % 2; }
}

// streaming layout (quant.cuh) -> tile layout; rows >= N read as zero bytes
__global__ void tile_repack_kernel(const uint8_t* __restrict__ w, int type, int64_t N, int K, uint8_t* __restrict__ out) {
    const int ks = K / QG_KSTEP, blob = qg_blob_bytes(type), ng = n_groups(K), gb = group_bytes(type);
    const int64_t n_tiles = (N + QG_ROWS - 1) / QG_ROWS, total = n_tiles * ks * (int64_t)blob;
    for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
        const int o = (int)(idx % blob); const int64_t u = idx / blob; const int kstep = (int)(u % ks); const int64_t tile = u / ks;
        int row, blk, i; tile_src(type, o, row, blk, i);
        const int64_t r = tile * QG_ROWS + row;
        uint8_t v = 0;
        if (r < N) {
            const int sb = type == T_Q8_0 ? kstep * 8 + blk : kstep;          // raw block index inside the row
            const int per_grp = type == T_Q8_0 ? 64 : 8, g = sb / per_grp;
            v = w[(((r >> 1) * ng + g) * 2 + (r & 1)) * (size_t)gb + dst_byte(type, sb % per_grp, i)];
        }
        out[idx] = v;
    }
}

// Per-role cycle trace of CTA 0 (build variant -DB200_QG_TRACE; tools/qgemm_trace.py): g_qgt[role][unit][stamp], clock64
#ifdef B200_QG_TRACE
__device__ long long g_qgt[6 * 64 * 8];
#define QGT(role, it, k) do { if (blockIdx.x == 0 && (it) < 64) g_qgt[((role) * 64 + (it)) * 8 + (k)] = clock64(); } while (0)
#else
#define QGT(role, it, k) do {} while (0)
#endif
struct QSeg { const uint8_t* tl; float* C; int type, n_tiles, n_rows, step0, tile0; };
struct QGemmParams {
    QSeg seg[3]; int n_seg, ks /*K-steps per tile*/, total_steps, n_tiles, B, ldc, mode, n_raw; float scale; float* ws; uint32_t slot_bytes;
    int whole;      // 1 = CTA ranges are aligned to tiles (no partials; for matrices with many more tiles than CTAs)
    int debug;      // profiling only (B200_QG_DEBUG; results are then wrong): 1 = skip the expansion math, 2 = skip the MMAs, 4 = no X loads,
                    // 8 = no weight loads, 16 = launch + TMEM alloc/dealloc only, 64 = skip the tensor-memory stores, 128 (with 1) = sleep instead of math
};

__device__ __forceinline__ int qg_seg_of_step(const QGemmParams& P, int s) { int i = 0; while (i + 1 < P.n_seg && s >= P.seg[i + 1].step0) i++; return i; }

// ------------------------------------------------------------------------------------ fp16x2 helpers
__device__ __forceinline__ uint32_t h2fma(uint32_t a, uint32_t b, uint32_t c) { uint32_t d; asm("fma.rn.f16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d; }
__device__ __forceinline__ uint32_t h2sub(uint32_t a, uint32_t b) { uint32_t d; asm("sub.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t h2mul(uint32_t a, uint32_t b) { uint32_t d; asm("mul.rn.f16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d; }
__device__ __forceinline__ uint32_t h2dup(float v) { uint32_t d; asm("cvt.rn.f16x2.f32 %0, %1, %1;" : "=r"(d) : "f"(v)); return d; }
__device__ __forceinline__ float h16f(uint32_t bits) { return __half2float(__ushort_as_half((unsigned short)bits)); }
// explicit shared-space loads (a generic pointer into the dynamic shared array compiles to LD.E, not LDS)
__device__ __forceinline__ uint4 lds128(uint32_t a) { uint4 v; asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint32_t lds16(uint32_t a) { unsigned short v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }

// (a & mask) | c in ONE LOP3: the mask is an immediate, c must sit in a register (an instruction has one immediate slot;
// written as C the compiler folds both constants and emits two LOP3s).  h2_magic() hides the constant from the folder.
template <uint32_t MASK> __device__ __forceinline__ uint32_t and_or_i(uint32_t a, uint32_t c) { uint32_t d; asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(a), "n"(MASK), "r"(c)); return d; }
__device__ __forceinline__ uint32_t h2_magic() { uint32_t m; asm volatile("mov.b32 %0, 0x64006400;" : "=r"(m)); return m; }
constexpr uint32_t H2_MAGIC = 0x64006400u /*1024*/, H2_M1024 = 0xE400E400u, H2_1152 = 0x64806480u, H2_1056 = 0x64206420u;

// ------------------------------------------------------------------------------------ tensor-memory stores / A-from-TMEM MMA
#define QG_R8(r, o) "r"(r[o]), "r"(r[o + 1]), "r"(r[o + 2]), "r"(r[o + 3]), "r"(r[o + 4]), "r"(r[o + 5]), "r"(r[o + 6]), "r"(r[o + 7])
// this warp's 32 lanes (TMEM lanes of its quadrant) x n consecutive 32-bit columns starting at taddr
__device__ __forceinline__ void tmem_st_x8(uint32_t taddr, const uint32_t* r) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"r"(taddr), QG_R8(r, 0) : "memory");
}
__device__ __forceinline__ void tmem_st_x32(uint32_t taddr, const uint32_t* r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
        "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};"
        ::"r"(taddr), QG_R8(r, 0), QG_R8(r, 8), QG_R8(r, 16), QG_R8(r, 24) : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem: 128 lanes x 8 columns = 128 x 16 fp16] . B[smem desc]^T
__device__ __forceinline__ void tcgen05_mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// ------------------------------------------------------------------------------------ per-type expansion, one lane = one row
// Each routine fills o[32] = the lane's 64 expanded elements as fp16 pairs.  A k-block (64 K positions) is 32 columns:
// column c holds positions (2c, 2c+1), i.e. elements (e, e+2) or (e+1, e+3) of an aligned group of four (kperm4).
// Q4_K / Q5_K: chunk j2 of the super-block (sub-blocks 2*j2: low nibbles -> columns 0..15, 2*j2+1: high -> 16..31)
template <bool Q5>
__device__ __forceinline__ void dq_k45(uint32_t raw, int row, int j2, uint32_t* o) {
    const uint4 hd = lds128(raw + row * 16);
    const uint32_t f = j2 == 0 ? hd.y : j2 == 1 ? __funnelshift_r(hd.y, hd.z, 24) : j2 == 2 ? __funnelshift_r(hd.z, hd.w, 16) : hd.w >> 8;
    const float d = h16f(hd.x & 0xFFFF) * QG_PRESCALE, dmin = h16f(hd.x >> 16) * QG_PRESCALE;
    const uint32_t S0 = h2dup(d * (float)(f & 63)), S1 = h2dup(d * (float)((f >> 6) & 63) * 0.0625f);
    const uint32_t M0 = h2dup(dmin * (float)((f >> 12) & 63)), M1 = h2dup(dmin * (float)((f >> 18) & 63));
    const uint32_t C0 = h2mul(S0, H2_M1024), C1 = h2mul(S1, H2_M1024), MG = h2_magic();
    const uint32_t qs = raw + 2048 + (Q5 ? 4096 : 0) + (j2 * 2) * 2048 + row * 16;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint4 q = lds128(qs + h * 2048);
        uint4 hq = make_uint4(0, 0, 0, 0);
        if (Q5) hq = lds128(raw + 2048 + h * 2048 + row * 16);
        const uint32_t xs[4] = {q.x, q.y, q.z, q.w}, hs[4] = {hq.x, hq.y, hq.z, hq.w};
#pragma unroll
        for (int w = 0; w < 4; w++) {                              // bytes l = 16h + 4w .. +3 -> columns 8h + 2w, +1 (low) and 16 + that (high)
            const uint32_t x = xs[w], y = x >> 8;
            uint32_t l02 = and_or_i<0x000F000Fu>(x, MG), h02 = and_or_i<0x00F000F0u>(x, MG), l13 = and_or_i<0x000F000Fu>(y, MG), h13 = and_or_i<0x00F000F0u>(y, MG);
            if (Q5) {
                const uint32_t a = hs[w] >> (2 * j2), b = a >> 8;
                l02 |= (a << 4) & 0x00100010u; h02 |= (a << 7) & 0x01000100u; l13 |= (b << 4) & 0x00100010u; h13 |= (b << 7) & 0x01000100u;
            }
            o[8 * h + 2 * w] = h2sub(h2fma(l02, S0, C0), M0); o[8 * h + 2 * w + 1] = h2sub(h2fma(l13, S0, C0), M0);
            o[16 + 8 * h + 2 * w] = h2sub(h2fma(h02, S1, C1), M1); o[16 + 8 * h + 2 * w + 1] = h2sub(h2fma(h13, S1, C1), M1);
        }
    }
}

// Q8_0: k-block kb of the K-step = blocks 2kb, 2kb+1 (columns 0..15, 16..31)
__device__ __forceinline__ void dq_q80(uint32_t raw, int row, int kb, uint32_t* o) {
    const uint32_t dd = lds32(raw + 32768 + (kb * 128 + row) * 4);
    const uint32_t D0 = h2dup(h16f(dd & 0xFFFF) * QG_PRESCALE), D1 = h2dup(h16f(dd >> 16) * QG_PRESCALE);
#pragma unroll
    for (int pc = 0; pc < 4; pc++) {
        const uint4 q = lds128(raw + ((kb * 4 + pc) * 128 + row) * 16);
        const uint32_t xs[4] = {q.x, q.y, q.z, q.w}, D = pc < 2 ? D0 : D1;
#pragma unroll
        for (int w = 0; w < 4; w++) {
            const uint32_t u = xs[w] ^ 0x80808080u;
            o[8 * pc + 2 * w] = h2mul(h2sub(__byte_perm(u, 0x64646464u, 0x4240), H2_1152), D);
            o[8 * pc + 2 * w + 1] = h2mul(h2sub(__byte_perm(u, 0x64646464u, 0x4341), H2_1152), D);
        }
    }
}

// Q6_K: for 128-half n and 16-run h, the 64 elements (grp 0..3) x (l in [16h, 16h+16)) of the row:
// o[0..7] = grp 0 -> k-block 2n columns 8h..8h+7; o[8..15] = grp 1 -> k-block 2n columns 16+8h..; o[16..23] = grp 2 and
// o[24..31] = grp 3 -> the same column ranges of k-block 2n+1
__device__ __forceinline__ void dq_q6k(uint32_t raw, int row, int n, int h, uint32_t* o) {
    const uint4 qa = lds128(raw + ((n * 4 + h) * 128 + row) * 16), qb = lds128(raw + ((n * 4 + 2 + h) * 128 + row) * 16);
    const uint4 qh = lds128(raw + 16384 + ((n * 2 + h) * 128 + row) * 16);
    const uint32_t sc = lds32(raw + 24576 + ((n * 2 + h) * 128 + row) * 4);
    const float d = h16f(lds16(raw + 26624 + row * 2)) * QG_PRESCALE;
    const uint32_t S0 = h2dup(d * (float)(int8_t)(sc & 0xFF)), S1 = h2dup(d * (float)(int8_t)((sc >> 8) & 0xFF)),
                   S2 = h2dup(d * (float)(int8_t)((sc >> 16) & 0xFF)), S3 = h2dup(d * (float)(int8_t)(sc >> 24));
    const uint32_t as[4] = {qa.x, qa.y, qa.z, qa.w}, bs[4] = {qb.x, qb.y, qb.z, qb.w}, hs[4] = {qh.x, qh.y, qh.z, qh.w};
    const uint32_t MG = h2_magic();
#define Q6V(lo, hi) and_or_i<0x00300030u>((hi), and_or_i<0x000F000Fu>((lo), MG))      /* (lo & 0x000F000F) | (hi & 0x00300030) | 1024: two LOP3 */
#pragma unroll
    for (int w = 0; w < 4; w++) {
        const uint32_t xa = as[w], xb = bs[w], hq = hs[w];
        o[2 * w]          = h2mul(h2sub(Q6V(xa, hq << 4), H2_1056), S0);
        o[2 * w + 1]      = h2mul(h2sub(Q6V(xa >> 8, hq >> 4), H2_1056), S0);
        o[8 + 2 * w]      = h2mul(h2sub(Q6V(xb, hq << 2), H2_1056), S1);
        o[8 + 2 * w + 1]  = h2mul(h2sub(Q6V(xb >> 8, hq >> 6), H2_1056), S1);
        o[16 + 2 * w]     = h2mul(h2sub(Q6V(xa >> 4, hq), H2_1056), S2);
        o[16 + 2 * w + 1] = h2mul(h2sub(Q6V(xa >> 12, hq >> 8), H2_1056), S2);
        o[24 + 2 * w]     = h2mul(h2sub(Q6V(xb >> 4, hq >> 2), H2_1056), S3);
        o[24 + 2 * w + 1] = h2mul(h2sub(Q6V(xb >> 12, hq >> 10), H2_1056), S3);
    }
#undef Q6V
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

// shared memory: X ring (2 K-steps, 1024-aligned) | raw-weight ring | barriers
struct QgSmem { uint8_t *x, *raw; uint64_t *raw_full, *raw_empty, *x_full, *a_full, *a_empty /* frees the A slot and the X slot of a K-step */, *acc_full, *acc_empty; uint32_t* tmem_slot; };
constexpr uint32_t QG_BAR_BYTES = 512;
__host__ __device__ inline size_t qg_smem_bytes(uint32_t slot_bytes, int n_raw) { return 1024 + QG_X_STEPS * QG_X_BYTES + (size_t)n_raw * slot_bytes + QG_BAR_BYTES; }

// ------------------------------------------------------------------------------------ the kernel
__global__ void __launch_bounds__(QG_THREADS, 1)
qgemm_kernel(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ QGemmParams P) {
    extern __shared__ uint8_t qg_smem_raw[];
    QgSmem S;
    {
        uint8_t* base = (uint8_t*)(((uintptr_t)qg_smem_raw + 1023) & ~(uintptr_t)1023);
        S.x = base; S.raw = S.x + QG_X_STEPS * QG_X_BYTES;
        uint64_t* b = (uint64_t*)(S.raw + (size_t)P.n_raw * P.slot_bytes);
        S.raw_full = b; S.raw_empty = b + QG_MAX_RAW; S.x_full = b + 2 * QG_MAX_RAW;
        S.a_full = S.x_full + QG_X_STEPS; S.a_empty = S.a_full + QG_A_STEPS; S.acc_full = S.a_empty + QG_A_STEPS; S.acc_empty = S.acc_full + 2;
        S.tmem_slot = (uint32_t*)(S.acc_empty + 2);
    }
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    pdl_launch_dependents();                           // the consuming row kernel may be scheduled; it waits for this grid to finish
    if (tid == 0) {
        for (int i = 0; i < P.n_raw; i++) { mbar_init(&S.raw_full[i], 1); mbar_init(&S.raw_empty[i], QG_DQ_WARPS / 2); }
        for (int i = 0; i < QG_X_STEPS; i++) mbar_init(&S.x_full[i], 1);
        for (int i = 0; i < QG_A_STEPS; i++) { mbar_init(&S.a_full[i], QG_DQ_WARPS / 2); mbar_init(&S.a_empty[i], 1); }
        for (int i = 0; i < 2; i++) { mbar_init(&S.acc_full[i], 1); mbar_init(&S.acc_empty[i], 4); }
        fence_barrier_init();
    }
    if (warp == 2) tcgen05_alloc(S.tmem_slot, QG_TMEM_COLS);
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem = *S.tmem_slot;

    const int G = gridDim.x, cta = blockIdx.x;
    const long long T = P.total_steps;
    int s_begin = qg_cta_first_step(cta, P.total_steps, G), s_end = qg_cta_first_step(cta + 1, P.total_steps, G);
    if (P.whole) { s_begin = (int)((long long)cta * P.n_tiles / G) * P.ks; s_end = (int)((long long)(cta + 1) * P.n_tiles / G) * P.ks; }
    if (P.debug & 16) s_end = s_begin;                 // profiling: launch + TMEM alloc/dealloc only

    if (warp == 0) {                                   // ===== weight producer: one bulk copy per (tile, K-step) blob =====
        if (lane == 0) {
            int it = 0;
            for (int s = s_begin; s < s_end; s++, it++) {
                const QSeg& sg = P.seg[qg_seg_of_step(P, s)];
                const int rs = it % P.n_raw; const uint32_t ph = (it / P.n_raw) & 1, bytes = qg_blob_bytes(sg.type);
                mbar_wait(&S.raw_empty[rs], ph ^ 1);
                QGT(0, it, 0);
                if (P.debug & 8) { mbar_arrive(&S.raw_full[rs]); continue; }
                mbar_expect_tx(&S.raw_full[rs], bytes);
                bulk_g2s(S.raw + (size_t)rs * P.slot_bytes, sg.tl + (size_t)(s - sg.step0) * bytes, bytes, &S.raw_full[rs]);
            }
        }
    } else if (warp == 1) {                            // ===== activation producer: four 64 x 64 fp16 boxes per K-step =====
        if (lane == 0) {
            pdl_wait();                                // the activations come from the preceding kernel (programmatic dependent launch)
            int it = 0;
            for (int s = s_begin; s < s_end; s++, it++) {
                const QSeg& sg = P.seg[qg_seg_of_step(P, s)];
                const int kstep = (s - sg.step0) % P.ks, xs = it % QG_X_STEPS;
                mbar_wait(&S.a_empty[xs], ((it / QG_X_STEPS) & 1) ^ 1);       // freed together with the A slot of the same K-step
                QGT(1, it, 0);
                if (P.debug & 4) { mbar_arrive(&S.x_full[xs]); continue; }
                mbar_expect_tx(&S.x_full[xs], QG_X_BYTES);
#pragma unroll
                for (int kb = 0; kb < 4; kb++) tma_load_2d(S.x + xs * QG_X_BYTES + kb * (QG_NB * 128), &tmap_x, kstep * QG_KSTEP + kb * 64, 0, &S.x_full[xs]);
            }
        }
    } else if (warp == 2) {                            // ===== MMA issuer: 16 MMAs (4 k-blocks x K=16) per K-step =====
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(QG_ROWS, QG_NB);
            int it = 0, piece = 0, s = s_begin;
            while (s < s_end) {
                const QSeg& sg = P.seg[qg_seg_of_step(P, s)];
                const int k0 = (s - sg.step0) % P.ks, n_st = min(P.ks - k0, s_end - s);
                const int buf = piece & 1;
                mbar_wait(&S.acc_empty[buf], ((piece >> 1) & 1) ^ 1);
                tcgen05_fence_after();
                for (int i = 0; i < n_st; i++, it++) {
                    const int as = it % QG_A_STEPS, xs = it % QG_X_STEPS;
                    QGT(2, it, 0);
                    mbar_wait(&S.a_full[as], (it / QG_A_STEPS) & 1);
                    QGT(2, it, 1);
                    mbar_wait(&S.x_full[xs], (it / QG_X_STEPS) & 1);
                    QGT(2, it, 2);
                    tcgen05_fence_after();
                    if (!(P.debug & 2)) {
#pragma unroll
                        for (int kb = 0; kb < 4; kb++) {
                            const uint64_t db = umma_desc_sw128(S.x + xs * QG_X_BYTES + kb * (QG_NB * 128));
#pragma unroll
                            for (int k = 0; k < 4; k++)
                                tcgen05_mma_f16_ts(tmem + buf * QG_NB, tmem + QG_TMEM_A0 + (as * 4 + kb) * 32 + k * 8, db + 2 * k, idesc, (i | kb | k) ? 1u : 0u);
                        }
                    }
                    tcgen05_commit(&S.a_empty[as]);                 // one commit per K-step frees both the A slot and the X slot (same ring depth)
                    QGT(2, it, 3);
                }
                tcgen05_commit(&S.acc_full[buf]);
                s += n_st; piece++;
            }
        }
    } else if (warp < QG_FIRST_EPI) {                  // ===== dequant warps: raw blob -> fp16 A operand in tensor memory =====
        // Two groups of 8 warps take alternate K-steps, so that one group computes while the other waits on its
        // barriers / tensor-memory stores (all warps in lockstep would expose every latency once per step).
        // In a group, warp (slab, half) expands rows 32*slab.. of portions 2*half and 2*half+1 of the step.
        const int idx = warp - QG_FIRST_DQ, grp = idx >> 3, half = (idx & 7) >> 2, row = (warp & 3) * 32 + lane;   // TMEM lanes of quadrant warp % 4 only
        const uint32_t lane_base = tmem + QG_TMEM_A0 + ((uint32_t)((warp & 3) * 32) << 16);
        for (int it = grp; s_begin + it < s_end; it += 2) {
            const int type = P.seg[qg_seg_of_step(P, s_begin + it)].type;
            const int rs = it % P.n_raw, as = it % QG_A_STEPS;
            const uint32_t raw = smem_u32(S.raw) + rs * P.slot_bytes;
            const uint32_t a0 = lane_base + as * 128;           // 4 k-blocks x 32 columns per K-step
            const int trole = (idx & 7) == 0 && lane == 0 ? 3 + grp : -1;
            if (trole >= 0) QGT(trole, it, 0);
            mbar_wait(&S.raw_full[rs], (it / P.n_raw) & 1);
            if (trole >= 0) QGT(trole, it, 1);
#pragma unroll
            for (int pp = 0; pp < 2; pp++) {
                const int p = 2 * half + pp;
                uint32_t o[32];
                if (P.debug & 1) {
#pragma unroll
                    for (int i = 0; i < 32; i++) o[i] = 0;
                    if (P.debug & 128) __nanosleep(150);
                } else if (type == T_Q8_0) dq_q80(raw, row, p, o);
                else if (type == T_Q6_K) dq_q6k(raw, row, p >> 1, p & 1, o);
                else if (type == T_Q4_K) dq_k45<false>(raw, row, p, o);
                else dq_k45<true>(raw, row, p, o);
                if (pp == 0) { if (trole >= 0) QGT(trole, it, 2); mbar_wait(&S.a_empty[as], ((it / QG_A_STEPS) & 1) ^ 1); tcgen05_fence_after(); if (trole >= 0) QGT(trole, it, 3); }
                else { __syncwarp(); if (lane == 0) mbar_arrive(&S.raw_empty[rs]); }     // every lane has its operands in registers
                if (type == T_Q6_K) {
                    const int n = p >> 1, h = p & 1;
                    tmem_st_x8(a0 + (2 * n) * 32 + 8 * h, o);          tmem_st_x8(a0 + (2 * n) * 32 + 16 + 8 * h, o + 8);
                    tmem_st_x8(a0 + (2 * n + 1) * 32 + 8 * h, o + 16); tmem_st_x8(a0 + (2 * n + 1) * 32 + 16 + 8 * h, o + 24);
                } else if (!(P.debug & 64)) tmem_st_x32(a0 + p * 32, o);
            }
            if (trole >= 0) QGT(trole, it, 4);
            tmem_st_wait();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&S.a_full[as]);
            if (trole >= 0) QGT(trole, it, 5);
        }
    } else {                                           // ===== epilogue warps: TMEM -> C (whole tile) or fp32 partial + last-arriver reduction =====
        const int q = warp & 3;                        // TMEM lane quadrant this warp may read
        pdl_wait();                                    // C / the partials buffer may still be read by the preceding kernel
        int piece = 0, s = s_begin;
        const float scale = P.scale * (1.f / QG_PRESCALE);
        while (s < s_end) {
            const int si = qg_seg_of_step(P, s); const QSeg& sg = P.seg[si];
            const int tl = (s - sg.step0) / P.ks, k0 = (s - sg.step0) % P.ks, n_st = min(P.ks - k0, s_end - s);
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
            float* dst = sg.C + n;
            if (n_st == P.ks) {
                if (n < sg.n_rows) {
#pragma unroll
                    for (int b = 0; b < QG_NB; b++)
                        if (b < P.B) { float v = __uint_as_float(r[b]) * scale; if (P.mode == 1) v += dst[(size_t)b * P.ldc]; dst[(size_t)b * P.ldc] = v; }
                }
            } else {
                // partial K range: park it (slot 0 = the first tile this CTA touches, slot 1 = its last); the row kernel that
                // consumes C (or qgemm_finish_kernel) sums the parts of the tile in CTA order — see QgSplit in prefill.cuh
                const int t0 = sg.step0 + tl * P.ks;                           // first global step of the tile
                float* part = P.ws + ((size_t)(cta * 2 + (s_begin >= t0 ? 0 : 1)) * QG_NB) * QG_ROWS + q * 32 + lane;
#pragma unroll
                for (int b = 0; b < QG_NB; b++) if (b < P.B) part[(size_t)b * QG_ROWS] = __uint_as_float(r[b]);
            }
            s += n_st; piece++;
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 2) tcgen05_dealloc(tmem, QG_TMEM_COLS);
}

// Completes C from the partials for callers that have no consuming row kernel (the test entry): one thread per 4 columns
__global__ void qgemm_finish_kernel(const __grid_constant__ QgSplit sp, float* __restrict__ C, int ldc, int B, int N) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < (long long)B * N; i += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(i / N), n = (int)(i % N);
        float v = C[(size_t)b * ldc + n]; qg_gather<1>(sp, b, n, &v); C[(size_t)b * ldc + n] = v;
    }
}

}  // namespace b200
