// decode.cuh — the persistent batch-1 decode kernel: embedding gather (K11), RMSNorm (K3), fused
// QKV GEMV + RoPE + KV-page write (K1/K4), split-K paged-KV attention (K5), wo + residual, fused
// gate/up GEMV + SiLU*mul (K6), down + residual, lm_head + argmax (K8 greedy).  One launch decodes
// n_steps tokens (forced tokens first = sequential prefill, then greedy).  Phases are separated by
// a grid barrier; the weight stream (stream.cuh Producer) is not.
#pragma once
#include "stream.cuh"

namespace b200 {

__device__ __forceinline__ void grid_barrier(unsigned* ctr, unsigned& target, int tid, int G) {
    consumer_sync();
    if (tid == 0) {
        target += (unsigned)G;
        __threadfence();
        red_release_add(ctr, 1u);
        while ((int)(ld_acquire_u32(ctr) - target) < 0) { }
        __threadfence();
    }
    consumer_sync();
}

// ------------------------------------------------------------------------------------ embedding
// Bit-exact dequant of one element of a row stored in the engine layout.
__device__ __forceinline__ float embed_elem(const uint8_t* w, int type, int n_grp, int row, int e) {
    const int gb = group_bytes(type), g = e / GROUP_ELEMS, eg = e % GROUP_ELEMS, be = block_elems(type);
    const uint8_t* base = w + (((size_t)(row >> 1) * n_grp + g) * 2 + (row & 1)) * gb;
    const int blk = eg / be;
    return dequant_elem(type, [&](int i) { return base[dst_byte(type, blk, i)]; }, eg % be);
}

// ------------------------------------------------------------------------------------ attention (bs1)
template <int VEC>   // head_dim = 32*VEC
__device__ void attn_item(const DecodeParams& P, const Smem& sm, int layer, int head, int split, int pos, int warp, int lane) {
    constexpr int HD = 32 * VEC;
    const int n_ctx = pos + 1, chunk = (n_ctx + P.n_splits - 1) / P.n_splits;
    const int p0 = split * chunk, p1 = min(n_ctx, p0 + chunk);
    const int kvh = head / (P.n_head / P.n_head_kv);
    float qv[VEC], acc[VEC];
#pragma unroll
    for (int i = 0; i < VEC; i++) { qv[i] = __ldcg(P.q + head * HD + lane * VEC + i) * P.kq_scale; acc[i] = 0.f; }
    float m = -INFINITY, l = 0.f;
    for (int p = p0 + warp; p < p1; p += CONSUMER_WARPS) {
        const __half* kp = P.kv.at(layer, 0, kvh, p) + lane * VEC;
        const __half* vp = P.kv.at(layer, 1, kvh, p) + lane * VEC;
        float kf[VEC], vf[VEC];
        if constexpr (VEC == 4) {
            uint2 kr = __ldcg((const uint2*)kp), vr = __ldcg((const uint2*)vp);
            float2 a = __half22float2(*(__half2*)&kr.x), b = __half22float2(*(__half2*)&kr.y);
            kf[0] = a.x; kf[1] = a.y; kf[2] = b.x; kf[3] = b.y;
            a = __half22float2(*(__half2*)&vr.x); b = __half22float2(*(__half2*)&vr.y);
            vf[0] = a.x; vf[1] = a.y; vf[2] = b.x; vf[3] = b.y;
        } else {
            uint32_t kr = __ldcg((const uint32_t*)kp), vr = __ldcg((const uint32_t*)vp);
            float2 a = __half22float2(*(__half2*)&kr); kf[0] = a.x; kf[1] = a.y;
            a = __half22float2(*(__half2*)&vr); vf[0] = a.x; vf[1] = a.y;
        }
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < VEC; i++) s += qv[i] * kf[i];
        s = warp_sum(s);
        const float mn = fmaxf(m, s), corr = __expf(m - mn), pe = __expf(s - mn);
        l = l * corr + pe;
#pragma unroll
        for (int i = 0; i < VEC; i++) acc[i] = acc[i] * corr + pe * vf[i];
        m = mn;
    }
    // cross-warp combine through shared memory (aliases the activation buffer, idle in this phase)
    float* scr = (float*)sm.xq;                     // [8 warps][HD + 2]
    consumer_sync();
#pragma unroll
    for (int i = 0; i < VEC; i++) scr[warp * (HD + 2) + lane * VEC + i] = acc[i];
    if (lane == 0) { scr[warp * (HD + 2) + HD] = m; scr[warp * (HD + 2) + HD + 1] = l; }
    consumer_sync();
    const int tid = warp * 32 + lane;
    if (tid < HD) {
        float M = -INFINITY;
#pragma unroll
        for (int w = 0; w < CONSUMER_WARPS; w++) M = fmaxf(M, scr[w * (HD + 2) + HD]);
        float a = 0.f, L = 0.f;
#pragma unroll
        for (int w = 0; w < CONSUMER_WARPS; w++) {
            const float mw = scr[w * (HD + 2) + HD];
            const float f = mw == -INFINITY ? 0.f : __expf(mw - M);
            a += f * scr[w * (HD + 2) + tid]; L += f * scr[w * (HD + 2) + HD + 1];
        }
        const size_t o = (size_t)head * P.n_splits + split;
        P.part_acc[o * HD + tid] = a;
        if (tid == 0) { P.part_ml[o * 2] = M; P.part_ml[o * 2 + 1] = L; }
    }
}

// ------------------------------------------------------------------------------------ epilogues
struct ArgBest { float v; int i; };

__device__ __forceinline__ void rope_pair(const DecodeParams& P, int d, int pos, float& a, float& b) {
    if (d < P.n_rot) {
        float sn, cs; sincosf((float)pos * __ldg(P.inv_freq + (d >> 1)), &sn, &cs);
        const float x = a, y = b; a = x * cs - y * sn; b = x * sn + y * cs;
    }
}

__device__ __forceinline__ void epilogue(const DecodeParams& P, const OpDesc& op, int pair, float a0, float a1, int pos, ArgBest& best, float* logits_dst) {
    const int row = 2 * pair, HD = P.head_dim;
    switch (op.epi) {
        case EPI_STORE: *(float2*)(op.out + row) = make_float2(a0 * op.scale, a1 * op.scale); break;
        case EPI_RESID: { float2 o = __ldcg((const float2*)(op.out + row)); o.x += op.scale * a0; o.y += op.scale * a1; *(float2*)(op.out + row) = o; } break;
        case EPI_SWIGLU: op.out[pair] = a0 / (1.f + __expf(-a0)) * a1; break;
        case EPI_LOGITS: {
            a0 *= op.scale; a1 *= op.scale; *(float2*)(logits_dst + row) = make_float2(a0, a1);
            if (a0 > best.v) { best.v = a0; best.i = row; }
            if (a1 > best.v) { best.v = a1; best.i = row + 1; }
        } break;
        case EPI_ROPE_Q: rope_pair(P, row % HD, pos, a0, a1); *(float2*)(op.out + row) = make_float2(a0, a1); break;
        case EPI_ROPE_K: rope_pair(P, row % HD, pos, a0, a1);
            *(__half2*)(P.kv.at(op.layer, 0, row / HD, pos) + row % HD) = __floats2half2_rn(a0, a1); break;
        case EPI_V: *(__half2*)(P.kv.at(op.layer, 1, row / HD, pos) + row % HD) = __floats2half2_rn(a0, a1); break;
    }
}

// ------------------------------------------------------------------------------------ one GEMV phase (consumers)
__device__ void consume_phase(const DecodeParams& P, const Phase& ph, const Smem& sm, RingCursor& cur, int c, int G,
                              int pos, int tid, ArgBest& best, float* logits_dst, long long* tr) {
    const int warp = tid >> 5, lane = tid & 31;
    // ---- activation prologue: build the int8 vector(s) in shared memory
    bool fam_std = false, fam_q6 = false;
    for (int o = 0; o < ph.n_ops; o++) { if (ph.op[o].type == T_Q6_K) fam_q6 = true; else fam_std = true; }
    const int n_grp = n_groups(ph.K);
    Smem s2 = sm;                                   // second layout family lives behind the first
    if (fam_std && fam_q6) { s2.xq += (size_t)n_grp * 2048; s2.xbs += n_grp * 128; s2.xd += n_grp * 8; }
    float rs = 1.f;
    if (ph.pro == PRO_NORM) rs = rms_scale(sm, ph.src, ph.K, P.rms_eps, tid);
    auto loader = [&](int u, float v[16]) {
        if (ph.pro == PRO_ATTN) {
            const int HD = P.head_dim, head = (16 * u) / HD, d0 = (16 * u) % HD, S = P.n_splits;
            float M = -INFINITY;
            for (int s = 0; s < S; s++) M = fmaxf(M, __ldcg(P.part_ml + ((size_t)head * S + s) * 2));
            float den = 0.f;
#pragma unroll
            for (int i = 0; i < 16; i++) v[i] = 0.f;
            for (int s = 0; s < S; s++) {
                const size_t o = (size_t)head * S + s;
                const float ms = __ldcg(P.part_ml + o * 2), f = ms == -INFINITY ? 0.f : __expf(ms - M);
                den += f * __ldcg(P.part_ml + o * 2 + 1);
                const float4* pa = (const float4*)(P.part_acc + o * HD + d0);
#pragma unroll
                for (int i = 0; i < 4; i++) { float4 t = __ldcg(pa + i); v[4 * i] += f * t.x; v[4 * i + 1] += f * t.y; v[4 * i + 2] += f * t.z; v[4 * i + 3] += f * t.w; }
            }
            const float inv = 1.f / den;
#pragma unroll
            for (int i = 0; i < 16; i++) v[i] *= inv;
        } else {
            load_run_f32(ph.src, u, v);
            if (ph.pro == PRO_NORM) {
                const float4* wv = (const float4*)(ph.norm_w + 16 * u);
#pragma unroll
                for (int i = 0; i < 4; i++) { float4 t = __ldg(wv + i); v[4 * i] *= rs * t.x; v[4 * i + 1] *= rs * t.y; v[4 * i + 2] *= rs * t.z; v[4 * i + 3] *= rs * t.w; }
            }
        }
    };
    if (fam_std && fam_q6) {
        // both layouts: quantise once per family (cheap: K = n_embd here)
        prologue_quantise(sm, T_Q4_K, ph.K, n_grp, tid, loader);
        prologue_quantise(s2, T_Q6_K, ph.K, n_grp, tid, loader);
    } else {
        prologue_quantise(sm, fam_q6 ? T_Q6_K : T_Q4_K, ph.K, n_grp, tid, loader);
    }

    if (tr && tid == 0) tr[1] = clock64();
    // ---- stream the CTA's slice
    int lo, hi; cta_slice(ph.total_pairs, c, G, lo, hi);
    int off = 0;
    for (int o = 0; o < ph.n_ops; o++) {
        const OpDesc& op = ph.op[o];
        const int a = max(lo, off) - off, b = min(hi, off + op.n_pairs) - off; off += op.n_pairs;
        if (a >= b) continue;
        const int gb = group_bytes(op.type); const uint32_t pu = 2u * gb;
        const Smem& sx = (op.type == T_Q6_K) ? s2 : sm;
        for (int rb = a; rb < b; rb += PAIRS_PER_STAGE) {
            const int np = min(PAIRS_PER_STAGE, b - rb);
            float acc0 = 0.f, acc1 = 0.f;
            for (int g = 0; g < op.n_grp; g++) {
                uint32_t gap; const uint32_t soff = cur.place(np * pu, sm.ring_bytes, gap);
                const uint32_t slot = cur.seq % NSLOT, parity = (cur.seq / NSLOT) & 1;
                mbar_wait(&sm.full[slot], parity);
                if (warp < np) {
                    XRegs x; load_x(sx, g, lane, x);
                    const uint8_t* unit = sm.ring + soff + warp * pu;
                    acc0 += lane_dot(op.type, unit, lane, x);
                    acc1 += lane_dot(op.type, unit + gb, lane, x);
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm.empty[slot]);
                cur.seq++;
            }
            if (warp < np) {
                acc0 = warp_sum(acc0); acc1 = warp_sum(acc1);
                if (lane == 0) epilogue(P, op, rb + warp, acc0, acc1, pos, best, logits_dst);
            }
        }
    }
}

// ------------------------------------------------------------------------------------ the kernel
struct DecodeRun {
    const int* forced; int n_forced;     // tokens fed regardless of argmax (prompt); forced[0] is the first input
    float* logits_all;                   // optional [n_steps][n_vocab]
};

__global__ void __launch_bounds__(CTA_THREADS, 2)
decode_kernel(const DecodeParams P, const DecodeRun R, const uint32_t ring_bytes, const int max_grp) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const Smem sm = carve_smem(smem_raw, ring_bytes, max_grp);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, c = blockIdx.x, G = gridDim.x;
    if (tid == 0) {
        for (int i = 0; i < NSLOT; i++) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], CONSUMER_WARPS); }
        fence_barrier_init();
    }
    __syncthreads();

    if (warp == CONSUMER_WARPS) {           // ===== producer warp =====
        Producer pr; pr.init();
        for (int step = 0; step < P.n_steps; step++) {
            const bool need_lm = (step + 1 >= R.n_forced) || R.logits_all;
            for (int i = 0; i < P.n_phases; i++) {
                const Phase& ph = P.phases[i];
                if (ph.kind != PH_GEMV) continue;
                if (ph.op[0].epi == EPI_LOGITS && !need_lm) continue;
                pr.run_phase(ph, sm, c, G, lane);
            }
        }
        pr.drain(sm);
        return;
    }

    // ===== consumer warps =====
    unsigned target = *(volatile unsigned*)P.epoch;
    const int pos0 = *(volatile int*)P.pos;
    int token = R.n_forced > 0 ? R.forced[0] : *(volatile int*)P.token;
    RingCursor cur; cur.seq = 0; cur.head = 0;
    for (int step = 0; step < P.n_steps; step++) {
        const int pos = pos0 + step;
        const bool need_lm = (step + 1 >= R.n_forced) || R.logits_all;
        float* logits_dst = R.logits_all ? R.logits_all + (size_t)step * P.n_vocab : P.logits;
        ArgBest best; best.v = -INFINITY; best.i = 0x7fffffff;
        for (int i = 0; i < P.n_phases; i++) {
            const Phase& ph = P.phases[i];
            long long* tr = nullptr;
            if (P.trace) { const int slot = c == 0 ? 0 : c == 1 ? 1 : c == G / 2 ? 2 : c == G - 1 ? 3 : -1; if (slot >= 0) tr = P.trace + (((size_t)step * P.n_phases + i) * 4 + slot) * 4; }
            if (tr && tid == 0) tr[0] = tr[1] = clock64();
            if (ph.kind == PH_EMBED) {
                const int ng = n_groups(P.n_embd);
                for (int e = c * CONSUMER_THREADS + tid; e < P.n_embd; e += G * CONSUMER_THREADS) {
                    float v = embed_elem(P.embd, P.embd_type, ng, token, e);
                    P.h[e] = P.emb_scale != 0.f ? v * P.emb_scale : v;
                }
            } else if (ph.kind == PH_ATTN) {
                for (int it = c; it < P.n_head * P.n_splits; it += G) {
                    if (P.head_dim == 128) attn_item<4>(P, sm, ph.layer, it / P.n_splits, it % P.n_splits, pos, warp, lane);
                    else attn_item<2>(P, sm, ph.layer, it / P.n_splits, it % P.n_splits, pos, warp, lane);
                }
            } else {
                if (ph.op[0].epi == EPI_LOGITS && !need_lm) continue;
                consume_phase(P, ph, sm, cur, c, G, pos, tid, best, logits_dst, tr);
                if (ph.op[0].epi == EPI_LOGITS) {          // publish this CTA's argmax candidate
                    consumer_sync();
                    if (lane == 0) { sm.red[2 * warp] = best.v; sm.red[2 * warp + 1] = __int_as_float(best.i); }
                    consumer_sync();
                    if (tid == 0) {
                        float bv = -INFINITY; int bi = 0x7fffffff;
                        for (int w = 0; w < CONSUMER_WARPS; w++) {
                            const float v = sm.red[2 * w]; const int ix = __float_as_int(sm.red[2 * w + 1]);
                            if (v > bv || (v == bv && ix < bi)) { bv = v; bi = ix; }
                        }
                        P.cand_val[c] = bv; P.cand_idx[c] = bi;
                    }
                }
            }
            if (tr && tid == 0) tr[2] = clock64();
            grid_barrier(P.bar_ctr, target, tid, G);
            if (tr && tid == 0) tr[3] = clock64();
        }
        // ---- next token: forced, or argmax over the per-CTA candidates (every CTA redundantly)
        int next;
        if (step + 1 < R.n_forced) next = R.forced[step + 1];
        else {
            float bv = -INFINITY; int bi = 0x7fffffff;
            for (int k = tid; k < G; k += CONSUMER_THREADS) {
                const float v = __ldcg(P.cand_val + k); const int ix = __ldcg(P.cand_idx + k);
                if (v > bv || (v == bv && ix < bi)) { bv = v; bi = ix; }
            }
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                const float v = __shfl_xor_sync(0xffffffffu, bv, o); const int ix = __shfl_xor_sync(0xffffffffu, bi, o);
                if (v > bv || (v == bv && ix < bi)) { bv = v; bi = ix; }
            }
            consumer_sync();
            if (lane == 0) { sm.red[2 * warp] = bv; sm.red[2 * warp + 1] = __int_as_float(bi); }
            consumer_sync();
            bv = -INFINITY; bi = 0x7fffffff;
            for (int w = 0; w < CONSUMER_WARPS; w++) {
                const float v = sm.red[2 * w]; const int ix = __float_as_int(sm.red[2 * w + 1]);
                if (v > bv || (v == bv && ix < bi)) { bv = v; bi = ix; }
            }
            next = bi;
        }
        if (c == 0 && tid == 0) P.out_tokens[step] = next;
        token = next;
    }
    if (c == 0 && tid == 0) { *P.pos = pos0 + P.n_steps; *P.token = token; *P.epoch = target; }
}

// ------------------------------------------------------------------------------------ repack (K9) and test kernels
// One thread per 16 output bytes; source bytes gathered through src_byte().
__global__ void repack_kernel(const uint8_t* __restrict__ raw, uint8_t* __restrict__ out, int type, long long N, int K,
                              long long row0, long long row_stride_pairs /*unused*/) {
    const int gb = group_bytes(type), ng = n_groups(K); const size_t rrb = raw_row_bytes(type, K);
    const long long chunks_per_row = (long long)ng * gb / 16, total = N * chunks_per_row;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / chunks_per_row; const int rem = (int)(i % chunks_per_row);
        const int g = rem / (gb / 16), o = (rem % (gb / 16)) * 16;
        uint8_t b[16];
#pragma unroll
        for (int k = 0; k < 16; k++) { long long s = src_byte(type, K, g, o + k); b[k] = s < 0 ? 0 : raw[(size_t)row * rrb + s]; }
        const long long orow = row0 + row;
        uint8_t* dst = out + ((((size_t)(orow >> 1)) * ng + g) * 2 + (orow & 1)) * gb + o;
        *(uint4*)dst = *(uint4*)b;
    }
}
// interleaving variant: raw row r goes to engine row 2r + parity (gate/up fusion)
__global__ void repack_interleave_kernel(const uint8_t* __restrict__ raw, uint8_t* __restrict__ out, int type, long long N, int K, int parity) {
    const int gb = group_bytes(type), ng = n_groups(K); const size_t rrb = raw_row_bytes(type, K);
    const long long chunks_per_row = (long long)ng * gb / 16, total = N * chunks_per_row;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / chunks_per_row; const int rem = (int)(i % chunks_per_row);
        const int g = rem / (gb / 16), o = (rem % (gb / 16)) * 16;
        uint8_t b[16];
#pragma unroll
        for (int k = 0; k < 16; k++) { long long s = src_byte(type, K, g, o + k); b[k] = s < 0 ? 0 : raw[(size_t)row * rrb + s]; }
        uint8_t* dst = out + (((size_t)row * ng + g) * 2 + parity) * gb + o;
        *(uint4*)dst = *(uint4*)b;
    }
}

// Dequantise a repacked matrix back to fp32 [N][K] (bit-exact check of layout + dequant).
__global__ void dequant_repacked_kernel(const uint8_t* __restrict__ w, int type, long long N, int K, float* __restrict__ out) {
    const int ng = n_groups(K);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < N * K; i += (long long)gridDim.x * blockDim.x)
        out[i] = embed_elem(w, type, ng, (int)(i / K), (int)(i % K));
}

}  // namespace b200
