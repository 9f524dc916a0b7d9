// quant.cuh — ggml block formats, the engine's HBM weight layout, repack (K9) and bit-exact dequant.
//
// HBM layout of one quantised matrix W[N x K] (N even, K % 256 == 0), chosen for streaming:
//   * K is cut into GROUPS of 2048 elements (8 super-blocks); a short last group is zero-padded.
//   * rows are taken in PAIRS (2p, 2p+1) — the pair shares the activation registers in the GEMV
//     and is the RoPE rotation pair / the (gate_i, up_i) SwiGLU pair.
//   * memory order is  [pair p][group g][row-in-pair r][group bytes]  so that the unit
//     (pair, group) is one contiguous, 16-byte-aligned run of 2*GB bytes: one bulk-async copy.
//   * inside a group the bytes are either raw ggml blocks (Q4_K, Q5_K: already 16 B friendly) or
//     split into lane-major planes (Q6_K, Q8_0: raw blocks are only 2-byte aligned) so that each
//     of the 32 lanes of a warp reads its 64 weights with conflict-free 128-bit shared loads.
// Same bytes as the GGUF file, rearranged; dequantising through this layout is bit-identical to
// gguf-py on the raw blocks (tests/test_gpu_dequant.py).  Block formats: gguf/quants.py:378-572,
// gguf/constants.py:4241-4246 (gguf-py 0.19.0).
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

namespace b200 {

enum GgmlType : int { T_F32 = 0, T_F16 = 1, T_Q8_0 = 8, T_Q4_K = 12, T_Q5_K = 13, T_Q6_K = 14, T_BF16 = 30 };

constexpr int GROUP_ELEMS = 2048;

__host__ __device__ inline int block_elems(int t) {
    switch (t) { case T_F32: case T_F16: case T_BF16: return 1; case T_Q8_0: return 32;
                 case T_Q4_K: case T_Q5_K: case T_Q6_K: return 256; default: return 0; }
}
__host__ __device__ inline int block_bytes(int t) {
    switch (t) { case T_F32: return 4; case T_F16: case T_BF16: return 2; case T_Q8_0: return 34;
                 case T_Q4_K: return 144; case T_Q5_K: return 176; case T_Q6_K: return 210; default: return 0; }
}
// bytes of one 2048-element group of one row in the engine layout
__host__ __device__ inline int group_bytes(int t) {
    switch (t) { case T_Q8_0: return 2176; case T_Q4_K: return 1152; case T_Q5_K: return 1408; case T_Q6_K: return 1680; default: return 0; }
}
__host__ __device__ inline bool is_streamable(int t) { return group_bytes(t) != 0; }
__host__ __device__ inline int n_groups(int K) { return (K + GROUP_ELEMS - 1) / GROUP_ELEMS; }
__host__ __device__ inline size_t repacked_bytes(int t, int64_t N, int K) {
    return (size_t)(N / 2) * n_groups(K) * 2 * group_bytes(t);
}
__host__ __device__ inline size_t raw_row_bytes(int t, int K) { return (size_t)K / block_elems(t) * block_bytes(t); }

// ---------------------------------------------------------------------------------------------
// Source-byte map: for output byte `o` (0 <= o < GB) of group g of a row, which raw row byte
// holds it (-1 = zero padding).  The repack kernel and the dequantiser both derive from the
// per-type maps below, so there is a single definition of each layout.
//
// Q4_K / Q5_K: raw blocks back to back.
// Q6_K group (1680 B): A-plane[32 lanes][16] | B-plane[32][16] | H-plane[32][16] | sc[32 lanes][4] | d[8] fp16
//    lane L = sb*4 + 2n + k  (sb: super-block in group, n: 128-half, k: 16-run)
//    A = ql[64n + 16k ..+16), B = ql[64n + 32 + 16k ..+16), H = qh[32n + 16k ..+16),
//    sc[L][g] = scales[8n + 2g + k]
// Q8_0 group (2176 B): plane p (0..3)[32 lanes][16] | d[64] fp16
//    lane L holds elements [64L, 64L+64): plane p = 16-run p of them; d[b] = scale of block b.
__host__ __device__ inline int64_t src_byte(int t, int K, int g, int o) {
    const int nsb_total = K / 256;
    if (t == T_Q4_K || t == T_Q5_K) {
        const int bb = block_bytes(t); int sb = g * 8 + o / bb;
        return sb < nsb_total ? (int64_t)sb * bb + o % bb : -1;
    }
    if (t == T_Q6_K) {
        int L, within, raw_off;
        if (o < 1536) {
            int plane = o / 512; L = (o % 512) / 16; within = o % 16;
            int n = (L % 4) / 2, k = L % 2;
            raw_off = plane == 0 ? 64 * n + 16 * k + within : plane == 1 ? 64 * n + 32 + 16 * k + within : 128 + 32 * n + 16 * k + within;
        } else if (o < 1664) {
            L = (o - 1536) / 4; int gq = (o - 1536) % 4; int n = (L % 4) / 2, k = L % 2;
            raw_off = 192 + 8 * n + 2 * gq + k;
        } else { L = ((o - 1664) / 2) * 4; raw_off = 208 + (o - 1664) % 2; }
        int sb = g * 8 + L / 4;
        return sb < nsb_total ? (int64_t)sb * 210 + raw_off : -1;
    }
    if (t == T_Q8_0) {
        int blk, raw_off;
        if (o < 2048) { int p = o / 512, L = (o % 512) / 16, within = o % 16; int e = L * 64 + p * 16 + within; blk = e / 32; raw_off = 2 + e % 32; }
        else { blk = (o - 2048) / 2; raw_off = (o - 2048) % 2; }
        int gb = g * 64 + blk;
        return gb < K / 32 ? (int64_t)gb * 34 + raw_off : -1;
    }
    return -1;
}

// Inverse map: raw byte i of block `blk` (block index inside the group: super-block for K-quants,
// 32-element block for Q8_0) -> byte offset inside the group.
__host__ __device__ inline int dst_byte(int t, int blk, int i) {
    if (t == T_Q4_K || t == T_Q5_K) return blk * block_bytes(t) + i;
    if (t == T_Q6_K) {
        if (i < 128) { int n = i / 64, r = i % 64, isB = r / 32, rr = r % 32, k = rr / 16, L = blk * 4 + 2 * n + k; return (isB ? 512 : 0) + L * 16 + rr % 16; }
        if (i < 192) { int ii = i - 128, n = ii / 32, k = (ii % 32) / 16, L = blk * 4 + 2 * n + k; return 1024 + L * 16 + ii % 16; }
        if (i < 208) { int s = i - 192, n = s / 8, gq = (s % 8) / 2, k = s % 2, L = blk * 4 + 2 * n + k; return 1536 + L * 4 + gq; }
        return 1664 + blk * 2 + (i - 208);
    }
    if (t == T_Q8_0) {
        if (i < 2) return 2048 + blk * 2 + i;
        int e = blk * 32 + (i - 2), L = e / 64, p = (e % 64) / 16; return p * 512 + L * 16 + e % 16;
    }
    return -1;
}

// ---------------------------------------------------------------------------------------------
// Activation layout companion: where the 16-run u (elements [16u,16u+16) of a 2048-group) of the
// int8 activation vector lives so that lane L finds its operands with 128-bit loads:
//   byte position  = plane*512 + L*16,   per-16 sum index = L*4 + plane
__host__ __device__ inline void act_slot(int t, int u, int& L, int& plane) {
    int sbl = u / 16, r = u % 16;
    if (t == T_Q6_K) { int n = r / 8, gq = (r % 8) / 2, k = r % 2; L = sbl * 4 + 2 * n + k; plane = gq; }
    else if (t == T_Q8_0) { L = u / 4; plane = u % 4; }
    else { L = sbl * 4 + r / 4; plane = r % 4; }      // Q4_K, Q5_K
}

// ---------------------------------------------------------------------------------------------
// Bit-exact dequantisation of element e of a raw block (float op order = gguf-py's).
__device__ __forceinline__ float h2f(uint16_t h) { return __half2float(__ushort_as_half(h)); }
__device__ __forceinline__ uint16_t ld16(const uint8_t* p) { return (uint16_t)p[0] | ((uint16_t)p[1] << 8); }

__device__ inline void scale_min_k4(int j, const uint8_t* q, int& sc, int& m) {   // quants.py:480-503
    if (j < 4) { sc = q[j] & 63; m = q[j + 4] & 63; }
    else { sc = (q[j + 4] & 0x0F) | ((q[j - 4] >> 6) << 4); m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

// ---------------------------------------------------------------------------------------------
// K-quant scale bytes, engine form.  ggml packs the sixteen 6-bit scale/min values of a super-block into 12 bytes
// in an order that costs ~25 ALU ops to undo per lane (quants.py:480-503).  The engine keeps the same 96 bits but
// orders them per lane: bytes 3j..3j+2 hold, for chunk j (sub-blocks 2j, 2j+1), the 24-bit field
//     sc[2j] | sc[2j+1] << 6 | m[2j] << 12 | m[2j+1] << 18
// so a lane needs one funnel shift and four bit-field extracts.  k4_reencode / k4_restore are exact inverses.
__host__ __device__ inline void k4_reencode(const uint8_t* q /*12 ggml bytes*/, uint8_t* out /*12 engine bytes*/) {
    for (int j = 0; j < 4; j++) {
        uint32_t f = 0;
        for (int k = 0; k < 2; k++) {
            const int s = 2 * j + k; uint32_t sc, m;
            if (s < 4) { sc = q[s] & 63; m = q[s + 4] & 63; }
            else { sc = (q[s + 4] & 0x0F) | ((q[s - 4] >> 6) << 4); m = (q[s + 4] >> 4) | ((q[s] >> 6) << 4); }
            f |= sc << (6 * k) | m << (12 + 6 * k);
        }
        out[3 * j] = (uint8_t)f; out[3 * j + 1] = (uint8_t)(f >> 8); out[3 * j + 2] = (uint8_t)(f >> 16);
    }
}
__host__ __device__ inline void k4_restore(const uint8_t* e /*12 engine bytes*/, uint8_t* q /*12 ggml bytes*/) {
    uint32_t sc[8], m[8];
    for (int j = 0; j < 4; j++) {
        const uint32_t f = (uint32_t)e[3 * j] | ((uint32_t)e[3 * j + 1] << 8) | ((uint32_t)e[3 * j + 2] << 16);
        sc[2 * j] = f & 63; sc[2 * j + 1] = (f >> 6) & 63; m[2 * j] = (f >> 12) & 63; m[2 * j + 1] = (f >> 18) & 63;
    }
    for (int i = 0; i < 4; i++) {
        q[i] = (uint8_t)((sc[i] & 63) | ((sc[i + 4] >> 4) << 6));
        q[i + 4] = (uint8_t)((m[i] & 63) | ((m[i + 4] >> 4) << 6));
        q[i + 8] = (uint8_t)((sc[i + 4] & 0x0F) | ((m[i + 4] & 0x0F) << 4));
    }
}

template <typename Rd>   // Rd: byte reader, rd(i) = i-th byte of the raw block
__device__ inline float dequant_elem(int t, Rd rd, int e) {
    if (t == T_Q8_0) {
        float d = h2f((uint16_t)rd(0) | ((uint16_t)rd(1) << 8));
        return __fmul_rn((float)(int8_t)rd(2 + e), d);
    }
    if (t == T_Q4_K || t == T_Q5_K) {
        float d = h2f((uint16_t)rd(0) | ((uint16_t)rd(1) << 8)), dmin = h2f((uint16_t)rd(2) | ((uint16_t)rd(3) << 8));
        uint8_t sb[12];
#pragma unroll
        for (int i = 0; i < 12; i++) sb[i] = rd(4 + i);
        int j = e / 32, l = e % 32, sc, m; scale_min_k4(j, sb, sc, m);
        float dl = __fmul_rn(d, (float)sc), ml = __fmul_rn(dmin, (float)m);
        int qoff = t == T_Q4_K ? 16 : 48;
        int v = (rd(qoff + 32 * (j / 2) + l) >> (4 * (j & 1))) & 0xF;
        if (t == T_Q5_K) v |= ((rd(16 + l) >> j) & 1) << 4;
        return __fsub_rn(__fmul_rn(dl, (float)v), ml);
    }
    if (t == T_Q6_K) {
        int half = e / 128, r = e % 128, grp = r / 32, l = r % 32;
        uint8_t qlb = rd(64 * half + 32 * (grp & 1) + l);
        int lo = grp < 2 ? (qlb & 0xF) : (qlb >> 4);
        int hi = (rd(128 + 32 * half + l) >> (2 * grp)) & 3;
        int q = (lo | (hi << 4)) - 32;
        float d = h2f((uint16_t)rd(208) | ((uint16_t)rd(209) << 8));
        float dl = __fmul_rn(d, (float)(int8_t)rd(192 + e / 16));
        return __fmul_rn(dl, (float)q);
    }
    return 0.f;
}

}  // namespace b200
