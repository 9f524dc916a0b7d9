// decode.cuh — the persistent batch-1 decode kernel: embedding gather (K11), RMSNorm (K3), fused
// QKV GEMV + RoPE + KV-page write (K1/K4), paged-KV attention (K5), wo + residual, fused gate/up
// GEMV + SiLU*mul (K6), down + residual, lm_head + argmax (K8 greedy).  One launch decodes n_steps
// tokens (forced tokens first = sequential prefill, then greedy).
//
// Synchronisation: there is NO grid barrier.  Every vector that crosses CTAs (hidden state, q,
// current K/V, attention output, SwiGLU activations, argmax candidates) is published as
// self-validating words — value and a generation tag in one atomic 4- or 8-byte store — and the
// consuming prologue polls the words it needs until their tags match ("LL" protocol).  A phase
// boundary therefore costs one store->load visibility latency instead of fence + atomic + spin +
// load, which matters because under a saturated weight stream every L2 round trip queues behind
// megabytes of bulk traffic (measured: ~4 us per round trip with a 27 MB backlog).
// Buffer-reuse safety (single buffers suffice) is argued in DESIGN.md §3.1.
#pragma once
#include <type_traits>

#include "stream.cuh"

namespace b200 {

// Per-warp cycle trace of one CTA (build variant -DB200_WTRACE, tools/wtrace_report.py): clock64 stamps per (step, phase, warp):
// 0 phase start, 1 own first poll done, 2 past the first prologue barrier, 3 own quantisation done, 4 prologue done,
// 5 first stage available, 6 stream loop done, 7 phase end.
#ifdef B200_WTRACE
__device__ long long* g_wtrace; __device__ int g_wtrace_cta;
__shared__ long long* s_wt[16];
#define WT(k) do { if ((threadIdx.x & 31) == 0) { long long* w_ = s_wt[threadIdx.x >> 5]; if (w_) w_[k] = clock64(); } } while (0)
#define WT_ONCE(k) do { if ((threadIdx.x & 31) == 0) { long long* w_ = s_wt[threadIdx.x >> 5]; if (w_ && !w_[k]) w_[k] = clock64(); } } while (0)
#else
#define WT(k) do {} while (0)
#define WT_ONCE(k) do {} while (0)
#endif

__device__ __forceinline__ long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return (long long)t; }

// ------------------------------------------------------------------------------------ tagged words
__device__ __forceinline__ uint32_t tag16_of(uint32_t seq) { return seq % 65535u + 1u; }            // never 0
__device__ __forceinline__ uint32_t ll16_pack(float v, uint32_t tag16) {
    v = fminf(fmaxf(v, -65504.f), 65504.f);
    return (tag16 << 16) | (uint32_t)__half_as_ushort(__float2half_rn(v));
}
__device__ __forceinline__ float ll16_val(uint32_t w) { return __half2float(__ushort_as_half((unsigned short)(w & 0xffffu))); }
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) { asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void st_relaxed_v2(void* p, uint32_t a, uint32_t b) { asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory"); }
__device__ __forceinline__ uint4 ld_relaxed_v4(const void* p) {
    uint4 r; asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory"); return r;
}
__device__ __forceinline__ uint2 ld_relaxed_v2(const void* p) {
    uint2 r; asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory"); return r;
}
// broadcast vectors exist in XREP copies: writers store all, reader CTA c polls copy c % XREP
__device__ __forceinline__ void st_rep_u32(uint32_t* base, int stride, int idx, uint32_t w) {
#pragma unroll
    for (int r = 0; r < XREP; r++) st_relaxed_u32(base + (size_t)r * stride + idx, w);
}
// 16 consecutive LL16 words (64 B), polled until every tag matches
__device__ __forceinline__ void ll16_poll_run(const uint32_t* buf, int u, uint32_t tag16, float v[16]) {
    const uint4* p = (const uint4*)(buf + 16 * u);
    const uint32_t want = tag16 << 16;
    uint4 r[4];
    for (;;) {
        uint32_t bad = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) { r[i] = ld_relaxed_v4(p + i); bad |= (r[i].x ^ want) | (r[i].y ^ want) | (r[i].z ^ want) | (r[i].w ^ want); }
        if ((bad >> 16) == 0) break;
        __nanosleep(40);
    }
#pragma unroll
    for (int i = 0; i < 4; i++) { v[4 * i] = ll16_val(r[i].x); v[4 * i + 1] = ll16_val(r[i].y); v[4 * i + 2] = ll16_val(r[i].z); v[4 * i + 3] = ll16_val(r[i].w); }
}
// two runs at once: all eight loads in flight before any tag is checked
__device__ __forceinline__ void ll16_poll_run2(const uint32_t* buf, int ua, int ub, uint32_t tag16, float va[16], float vb[16]) {
    const uint4* pa = (const uint4*)(buf + 16 * ua); const uint4* pb = (const uint4*)(buf + 16 * ub);
    const uint32_t want = tag16 << 16;
    uint4 r[8];
    for (;;) {
        uint32_t bad = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) { r[i] = ld_relaxed_v4(pa + i); r[4 + i] = ld_relaxed_v4(pb + i); }
#pragma unroll
        for (int i = 0; i < 8; i++) bad |= (r[i].x ^ want) | (r[i].y ^ want) | (r[i].z ^ want) | (r[i].w ^ want);
        if ((bad >> 16) == 0) break;
        __nanosleep(40);
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        va[4 * i] = ll16_val(r[i].x); va[4 * i + 1] = ll16_val(r[i].y); va[4 * i + 2] = ll16_val(r[i].z); va[4 * i + 3] = ll16_val(r[i].w);
        vb[4 * i] = ll16_val(r[4 + i].x); vb[4 * i + 1] = ll16_val(r[4 + i].y); vb[4 * i + 2] = ll16_val(r[4 + i].z); vb[4 * i + 3] = ll16_val(r[4 + i].w);
    }
}
// N consecutive LL32 elements ({bits, tag32} pairs), N even, polled
template <int N>
__device__ __forceinline__ void ll32_poll(const uint2* buf, uint32_t tag, float* out) {
    uint4 r[N / 2];
    for (;;) {
        uint32_t bad = 0;
#pragma unroll
        for (int i = 0; i < N / 2; i++) { r[i] = ld_relaxed_v4((const uint4*)buf + i); bad |= (r[i].y ^ tag) | (r[i].w ^ tag); }
        if (!bad) break;
    }
#pragma unroll
    for (int i = 0; i < N / 2; i++) { out[2 * i] = __uint_as_float(r[i].x); out[2 * i + 1] = __uint_as_float(r[i].z); }
}

// ------------------------------------------------------------------------------------ embedding
// Bit-exact dequant of one element of a row stored in the engine layout.
__device__ __forceinline__ float embed_elem(const uint8_t* w, int type, int n_grp, int row, int e) {
    const int gb = group_bytes(type), g = e / GROUP_ELEMS, eg = e % GROUP_ELEMS, be = block_elems(type);
    const uint8_t* base = w + (((size_t)(row >> 1) * n_grp + g) * 2 + (row & 1)) * gb;
    const int blk = eg / be;
    if (type == T_Q4_K || type == T_Q5_K) {      // scale bytes are stored re-encoded (quant.cuh k4_reencode)
        uint8_t q[12]; k4_restore(base + dst_byte(type, blk, 4), q);
        return dequant_elem(type, [&](int i) { return (i >= 4 && i < 16) ? q[i - 4] : base[dst_byte(type, blk, i)]; }, eg % be);
    }
    return dequant_elem(type, [&](int i) { return base[dst_byte(type, blk, i)]; }, eg % be);
}

// ------------------------------------------------------------------------------------ attention (bs1)
// One item = (head, split).  Each warp owns a strided set of positions and keeps NB of them in flight at once; the
// cached K/V loads are issued BEFORE q is polled (they do not depend on this token), so the exchange wait and the
// first memory round trip overlap.
template <int VEC, bool TR>   // head_dim = 32*VEC
__device__ void attn_item(const DecodeParams& P, const Smem& sm, int layer, int head, int split, int pos, uint32_t seq_in, uint32_t seq_out, int warp, int lane, long long* tr) {
    if constexpr (!TR) tr = nullptr;                                   // variants without the trace stamps: every `if (tr ...)` folds away
    constexpr int HD = 32 * VEC, NB = 10;     // 14 warps x 10 positions: a 128-position split (context 512 over 4 splits) is one batch
    typedef typename std::conditional<VEC == 4, uint2, uint32_t>::type raw_t;     // VEC fp16 values
    const int n_ctx = pos + 1, chunk = (n_ctx + P.n_splits - 1) / P.n_splits;
    const int p0 = split * chunk, p1 = min(n_ctx, p0 + chunk);
    const int kvh = head / (P.n_head / P.n_head_kv), KVD = P.n_head_kv * HD;
    auto unpack = [](const raw_t& r, float* f) {
        if constexpr (VEC == 4) { float2 a = __half22float2(*(const __half2*)&r.x), b = __half22float2(*(const __half2*)&r.y); f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; }
        else { float2 a = __half22float2(*(const __half2*)&r); f[0] = a.x; f[1] = a.y; }
    };
    auto load_batch = [&](int pb, raw_t* kr, raw_t* vr) {               // cached positions only; the current one is filled in later
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const int p = pb + b * CONSUMER_WARPS;
            if (p < p1 && p < pos) {
                const int page = sm.btab[p >> 4];                     // block table cached in shared memory: no dependent L2 load
                kr[b] = __ldcg((const raw_t*)(P.kv.at_page(page, layer, 0, kvh, p) + lane * VEC));
                vr[b] = __ldcg((const raw_t*)(P.kv.at_page(page, layer, 1, kvh, p) + lane * VEC));
            }
        }
    };
    raw_t kr[NB], vr[NB];
    load_batch(p0 + warp, kr, vr);
    float qv[VEC], acc[VEC];
    {   // q of this head: LL32, tag of the QKV phase
        float t[VEC];
        ll32_poll<VEC>(P.q_ll + head * HD + lane * VEC, seq_in, t);
        __syncwarp();                                                  // lanes leave the poll at different times; shuffles follow
#pragma unroll
        for (int i = 0; i < VEC; i++) { qv[i] = t[i] * P.kq_scale; acc[i] = 0.f; }
    }
    if (warp == 0 && lane == 0) sm.ctrl[0] = 0x7fffffff;             // q has arrived: the producer may run ahead again
    if (tr && warp == 0 && lane == 0) tr[1] = gtime();
    // A batch is processed as a block: all NB dot products, then their NB warp reductions interleaved (independent shuffle
    // chains instead of NB dependent ones), then one online-softmax update for the block.
    float m = -INFINITY, l = 0.f;
    for (int pb = p0 + warp; pb < p1; pb += CONSUMER_WARPS * NB) {
        if (pb != p0 + warp) load_batch(pb, kr, vr);
        float sc[NB];
#pragma unroll
        for (int b = 0; b < NB; b++) {
            const int p = pb + b * CONSUMER_WARPS;
            if (p == pos && p < p1) {                                     // the token being decoded: tagged exchange words -> the same raw form
                const uint32_t want = tag16_of(seq_in) << 16;
                const uint32_t* kw = P.kv_ll + kvh * HD + lane * VEC; const uint32_t* vw = P.kv_ll + KVD + kvh * HD + lane * VEC;
                uint32_t kk[VEC], vv[VEC];
                for (;;) {
                    uint32_t bad = 0;
                    if constexpr (VEC == 4) { uint4 a = ld_relaxed_v4(kw), c2 = ld_relaxed_v4(vw); kk[0] = a.x; kk[1] = a.y; kk[2] = a.z; kk[3] = a.w; vv[0] = c2.x; vv[1] = c2.y; vv[2] = c2.z; vv[3] = c2.w; }
                    else { uint2 a = ld_relaxed_v2(kw), c2 = ld_relaxed_v2(vw); kk[0] = a.x; kk[1] = a.y; vv[0] = c2.x; vv[1] = c2.y; }
#pragma unroll
                    for (int i = 0; i < VEC; i++) bad |= (kk[i] ^ want) | (vv[i] ^ want);
                    if ((bad >> 16) == 0) break;
                }
                if constexpr (VEC == 4) {
                    kr[b] = make_uint2((kk[0] & 0xffffu) | (kk[1] << 16), (kk[2] & 0xffffu) | (kk[3] << 16));
                    vr[b] = make_uint2((vv[0] & 0xffffu) | (vv[1] << 16), (vv[2] & 0xffffu) | (vv[3] << 16));
                } else { kr[b] = (kk[0] & 0xffffu) | (kk[1] << 16); vr[b] = (vv[0] & 0xffffu) | (vv[1] << 16); }
            }
            float d = 0.f;
            if (p < p1) {                                                 // warp-uniform
                float kf[VEC]; unpack(kr[b], kf);
#pragma unroll
                for (int i = 0; i < VEC; i++) d += qv[i] * kf[i];
            }
            sc[b] = d;
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
#pragma unroll
            for (int b = 0; b < NB; b++) sc[b] += __shfl_xor_sync(0xffffffffu, sc[b], o);
        }
        float mn = m;
#pragma unroll
        for (int b = 0; b < NB; b++) { if (pb + b * CONSUMER_WARPS >= p1) sc[b] = -INFINITY; mn = fmaxf(mn, sc[b]); }
        const float corr = m == -INFINITY ? 0.f : __expf(m - mn);        // mn is finite: the batch holds at least one position
        l *= corr;
#pragma unroll
        for (int i = 0; i < VEC; i++) acc[i] *= corr;
#pragma unroll
        for (int b = 0; b < NB; b++) if (pb + b * CONSUMER_WARPS < p1) {
            const float pe = __expf(sc[b] - mn);
            float vf[VEC]; unpack(vr[b], vf);
            l += pe;
#pragma unroll
            for (int i = 0; i < VEC; i++) acc[i] += pe * vf[i];
        }
        m = mn;
    }
    if (tr && warp == 0 && lane == 0) tr[2] = gtime();
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
        if (P.n_splits == 1) st_rep_u32(P.att_ll, P.att_stride, head * HD + tid, ll16_pack(a / L, tag16_of(seq_out)));
        else if (split != 0) {
            const size_t o = ((size_t)head * P.n_splits + split) * (HD + 2);
            st_relaxed_v2(P.part_ll + o + tid, __float_as_uint(a), seq_out);
            if (tid == 0) { st_relaxed_v2(P.part_ll + o + HD, __float_as_uint(M), seq_out); st_relaxed_v2(P.part_ll + o + HD + 1, __float_as_uint(L), seq_out); }
        } else { scr[CONSUMER_WARPS * (HD + 2) + tid] = a; if (tid == 0) { scr[CONSUMER_WARPS * (HD + 2) + HD] = M; scr[CONSUMER_WARPS * (HD + 2) + HD + 1] = L; } }
    }
    if (tr && warp == 0 && lane == 0) tr[3] = tr[4] = tr[5] = gtime();
    if (P.n_splits > 1 && split == 0) {
        // This CTA is the head's combiner: gather the other splits' partials (all loads in flight together,
        // re-polled until every tag matches), merge with its own, publish the head's output once.
        const int S = P.n_splits;
        float* own = scr + CONSUMER_WARPS * (HD + 2);            // [HD + 2]: a[], M, L of split 0
        float* mls = own + HD + 2;                               // [16][2]
        if (tid >= HD && tid < HD + S - 1) {
            const int s = tid - HD + 1; float t[2];
            ll32_poll<2>(P.part_ll + ((size_t)head * S + s) * (HD + 2) + HD, seq_out, t);
            mls[2 * s] = t[0]; mls[2 * s + 1] = t[1];
        }
        uint32_t av[16];
        if (tid < HD) {
            for (;;) {
                uint32_t bad = 0;
#pragma unroll
                for (int s = 1; s < 16; s++) if (s < S) { uint2 w = ld_relaxed_v2(P.part_ll + ((size_t)head * S + s) * (HD + 2) + tid); av[s] = w.x; bad |= w.y ^ seq_out; }
                if (!bad) break;
            }
        }
        consumer_sync();
        if (tr && warp == 0 && lane == 0) tr[4] = gtime();
        if (tid < HD) {
            float M = own[HD];
            for (int s = 1; s < S; s++) M = fmaxf(M, mls[2 * s]);
            const float f0 = own[HD] == -INFINITY ? 0.f : __expf(own[HD] - M);
            float a = f0 * own[tid], L = f0 * own[HD + 1];
#pragma unroll
            for (int s = 1; s < 16; s++) if (s < S) {
                const float ms = mls[2 * s], f = ms == -INFINITY ? 0.f : __expf(ms - M);
                a += f * __uint_as_float(av[s]); L += f * mls[2 * s + 1];
            }
            st_rep_u32(P.att_ll, P.att_stride, head * HD + tid, ll16_pack(a / L, tag16_of(seq_out)));
        }
        if (tr && warp == 0 && lane == 0) tr[5] = gtime();
    }
}

// ------------------------------------------------------------------------------------ tensor-parallel exchange (K10)
// All-reduce of a row pair's partial sums over the ranks, fused into the residual epilogue: every rank stores its two
// fp32 partials as tagged words straight into every peer's receive buffer (NVLink peer stores), then polls its OWN
// buffer for the peers' words and adds the tp_world partials in rank order — the same order on every rank, so the
// residual stream stays bit-identical across GPUs.  No extra launch, no barrier: one store -> poll latency per
// row-parallel product.  The buffer has two halves (wo | down): a half is rewritten one layer later, which a peer can
// only reach after this rank has sent — hence already read — the current values.
__device__ __forceinline__ void st_sys_v2(void* p, uint32_t a, uint32_t b) { asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory"); }
__device__ __forceinline__ uint2 ld_sys_v2(const void* p) {
    uint2 r; asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p) : "memory"); return r;
}
constexpr unsigned TP_SPIN_BUDGET = 1u << 26;            // polls before a peer is declared lost (~30 s: the ranks' hosts launch the same step at slightly different times)

// whole warp; a0/a1 hold the same values in every lane on entry and on exit
__device__ __forceinline__ void tp_allreduce_pair(const DecodeParams& P, int which, int row, uint32_t seq, int lane, float& a0, float& a1) {
    const int W = P.tp_world, me = P.tp_rank, D = P.n_embd;
    const int k = lane >> 1, comp = lane & 1;                         // lane -> (k-th other rank, component)
    const int other = k + (k >= me ? 1 : 0);
    float got = 0.f;
    if (k < W - 1) {
        st_sys_v2(P.rx_peer[other] + ((size_t)(which * W + me) * D + row + comp), __float_as_uint(comp ? a1 : a0), seq);
        const uint2* src = P.rx_peer[me] + ((size_t)(which * W + other) * D + row + comp);
        unsigned spins = *(volatile unsigned*)P.tp_error ? TP_SPIN_BUDGET : 0u;      // a peer was lost earlier in this launch: do not wait again
        for (;;) {
            const uint2 w = ld_sys_v2(src);
            if (w.y == seq) { got = __uint_as_float(w.x); break; }
            if (++spins > TP_SPIN_BUDGET) { *P.tp_error = 1u; break; }
        }
    }
    __syncwarp();                                                     // lanes leave the poll at different times (and half of them skip it)
    float s0 = 0.f, s1 = 0.f;
    for (int r = 0; r < W; r++) {                                     // rank order, identical everywhere
        const int kk = r - (r > me ? 1 : 0);
        const float p0 = __shfl_sync(0xffffffffu, got, 2 * kk), p1 = __shfl_sync(0xffffffffu, got, 2 * kk + 1);
        s0 += r == me ? a0 : p0; s1 += r == me ? a1 : p1;
    }
    a0 = s0; a1 = s1;
}

// ------------------------------------------------------------------------------------ epilogues
struct ArgBest { float v; int i; };

__device__ __forceinline__ void rope_pair(const DecodeParams& P, const Smem& sm, int d, float& a, float& b) {
    if (d < P.n_rot) {
        const float2 cs = sm.rope[d >> 1];           // filled once per token (decode_kernel), off the critical path
        const float x = a, y = b; a = x * cs.x - y * cs.y; b = x * cs.y + y * cs.x;
    }
}

// lp: index of the pair inside this CTA's slice of the op (for the owned hidden-state rows)
__device__ __forceinline__ void epilogue(const DecodeParams& P, const OpDesc& op, const Smem& sm, int pair, int lp, float a0, float a1, int pos,
                                         uint32_t seq, ArgBest& best, float* logits_dst) {
    const int row = 2 * pair, HD = P.head_dim; const uint32_t t16 = tag16_of(seq);
    switch (op.epi) {
        case EPI_STORE: *(float2*)(op.out + row) = make_float2(a0 * op.scale, a1 * op.scale); break;
        case EPI_RESID: {
            const float h0 = sm.hown[2 * lp] + op.scale * a0, h1 = sm.hown[2 * lp + 1] + op.scale * a1;
            sm.hown[2 * lp] = h0; sm.hown[2 * lp + 1] = h1;
            const uint32_t w0 = ll16_pack(h0, t16), w1 = ll16_pack(h1, t16);
#pragma unroll
            for (int r = 0; r < XREP; r++) st_relaxed_v2(P.h_ll + (size_t)r * P.h_stride + row, w0, w1);
        } break;
        case EPI_SWIGLU: st_rep_u32(P.act_ll, P.act_stride, pair, ll16_pack(a0 / (1.f + __expf(-a0)) * a1, t16)); break;
        case EPI_LOGITS: {
            a0 *= op.scale; a1 *= op.scale;
            logits_dst[row] = a0; if (row + 1 < op.n_rows) logits_dst[row + 1] = a1;
            if (a0 > best.v) { best.v = a0; best.i = P.vocab_base + row; }
            if (row + 1 < op.n_rows && a1 > best.v) { best.v = a1; best.i = P.vocab_base + row + 1; }
        } break;
        case EPI_ROPE_Q:
            rope_pair(P, sm, row % HD, a0, a1);
            st_relaxed_v2(P.q_ll + row, __float_as_uint(a0), seq); st_relaxed_v2(P.q_ll + row + 1, __float_as_uint(a1), seq); break;
        case EPI_ROPE_K: {
            rope_pair(P, sm, row % HD, a0, a1);
            const __half2 hv = __floats2half2_rn(a0, a1);
            *(__half2*)(P.kv.at_page(sm.btab[pos >> 4], op.layer, 0, row / HD, pos) + row % HD) = hv;       // block table from shared memory: no dependent global load
            st_relaxed_v2(P.kv_ll + row, (t16 << 16) | __half_as_ushort(__low2half(hv)), (t16 << 16) | __half_as_ushort(__high2half(hv)));
        } break;
        case EPI_V: {
            const __half2 hv = __floats2half2_rn(a0, a1);
            *(__half2*)(P.kv.at_page(sm.btab[pos >> 4], op.layer, 1, row / HD, pos) + row % HD) = hv;
            st_relaxed_v2(P.kv_ll + P.n_head_kv * HD + row, (t16 << 16) | __half_as_ushort(__low2half(hv)), (t16 << 16) | __half_as_ushort(__high2half(hv)));
        } break;
    }
}

// ------------------------------------------------------------------------------------ streaming one op's slice (consumers)
// Pairs [a, b) of `op`, eight pairs per stage (one per warp of a group).  A warp visits only the stages of its group's
// row blocks: their sequence numbers follow from the slice's stage order (seg_stage_index), the ring offset of a stage
// is published by the producer next to its barrier.  seq0: sequence number of the slice's first stage.
template <int TYPE, bool XHOLD, bool TP, bool TR>
__device__ __forceinline__ void stream_op(const DecodeParams& P, const OpDesc& op, const Smem& sm, const Smem& sx, uint32_t seq0, int a, int b,
                                          int pos, int warp, int lane, uint32_t seq, float in_scale, ArgBest& best, float* logits_dst, long long* tr) {
    if constexpr (!TR) tr = nullptr;
    constexpr int GB = TYPE == T_Q4_K ? 1152 : TYPE == T_Q5_K ? 1408 : TYPE == T_Q6_K ? 1680 : 2176;
    constexpr uint32_t PU = 2u * GB;
    constexpr int SG = stage_groups(TYPE);
    const int n_grp = op.n_grp, spr = (n_grp + SG - 1) / SG;
    const int wg = warp / GROUP_WARPS, wl = warp % GROUP_WARPS;
    const int n_rb = (b - a + PAIRS_PER_STAGE - 1) / PAIRS_PER_STAGE;
    // XHOLD (inputs of at most SG groups, i.e. one stage per row block): the warp's slice of the activation vector stays in
    // registers for all of its row blocks instead of being re-read from shared memory for every stage — in a phase with many
    // row blocks per warp (gate/up, lm_head) those reads were 40 % of the shared-memory wavefronts
    XRegs xh0, xh1;                                                     // XHOLD: exactly two groups (callers guarantee n_grp == SG == 2)
    if constexpr (XHOLD) { load_x(sx, 0, lane, xh0); load_x(sx, 1, lane, xh1); }
    for (int rbi = wg; rbi < n_rb; rbi += WARP_GROUPS) {
        const int rb = a + rbi * PAIRS_PER_STAGE, np = min(PAIRS_PER_STAGE, b - rb);
        const bool mine = wl < np;
        const int pp = rbi >> 1, nb = min(2, n_rb - 2 * pp);
        uint32_t sq = seq0 + pp * 2 * spr + (rbi & 1);                    // seg_stage_index(rbi, 0, ...), then += nb per stage
        float acc0 = 0.f, acc1 = 0.f;
        for (int gi = 0; gi < spr; gi++, sq += nb) {
            const int g0 = gi * SG, ng = min(SG, n_grp - g0); const uint32_t cb = ng * PU;
            const uint32_t slot = sq % NSLOT;
            long long w0 = 0; if (tr && warp == 0 && lane == 0) w0 = clock64();
            mbar_wait(&sm.full[slot], (sq / NSLOT) & 1);
            if (tr && warp == 0 && lane == 0) { tr[7] += clock64() - w0; if (rbi == 0 && gi == 0) tr[4] = gtime(); }   // tr[7]: cycles warp 0 waited for data
            WT_ONCE(5);
            if (mine) {
                const uint8_t* unit = sm.ring + sm.q_off[slot] + wl * cb;
                if constexpr (XHOLD) {
                    acc0 += lane_dot<TYPE>(unit, lane, xh0); acc1 += lane_dot<TYPE>(unit + GB, lane, xh0);
                    acc0 += lane_dot<TYPE>(unit + PU, lane, xh1); acc1 += lane_dot<TYPE>(unit + PU + GB, lane, xh1);
                } else {
#pragma unroll
                    for (int gg = 0; gg < SG; gg++) if (gg < ng) {
                        XRegs x; load_x(sx, g0 + gg, lane, x);
                        acc0 += lane_dot<TYPE>(unit + gg * PU, lane, x);
                        acc1 += lane_dot<TYPE>(unit + gg * PU + GB, lane, x);
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.empty[slot]);
        }
        if (tr && warp == 0 && lane == 0) tr[5] = gtime();
        WT(6);
        if (mine) {
            acc0 = warp_sum(acc0); acc1 = warp_sum(acc1);
            if constexpr (TP) { if (P.tp_world > 1 && op.epi == EPI_RESID) tp_allreduce_pair(P, op.tp_half, 2 * (rb + wl), seq, lane, acc0, acc1); }
            if (lane == 0) epilogue(P, op, sm, rb + wl, rb + wl - a, acc0 * in_scale, acc1 * in_scale, pos, seq, best, logits_dst);
        }
    }
}

// ------------------------------------------------------------------------------------ one GEMV phase (consumers)
// seq: this phase's generation number; inputs carry seq-1's tags.
template <int CFG>
__device__ void consume_phase(const DecodeParams& P, const Phase& ph, const Smem& sm, uint32_t& stage_seq, int c, int G,
                              int pos, int tid, uint32_t seq, ArgBest& best, float* logits_dst, long long* tr) {
    if constexpr ((CFG & DK_TRACE) == 0) tr = nullptr;
    const int warp = tid >> 5, lane = tid & 31;
    // ---- activation prologue: build the int8 vector(s) in shared memory
    bool fam_std = false, fam_q6 = false;
    for (int o = 0; o < ph.n_ops; o++) { if (ph.op[o].type == T_Q6_K) fam_q6 = true; else fam_std = true; }
    const int n_grp = n_groups(ph.K);
    Smem s2 = sm;                                   // second layout family lives behind the first
    if (fam_std && fam_q6) { s2.xq += (size_t)n_grp * 2048; s2.xbs += n_grp * 128; s2.xd += n_grp * 8; }
    const uint32_t tin16 = tag16_of(seq - 1);
    const int n_runs = ph.K / 16, n_runs_pad = n_grp * 128;
    const int rep = c % XREP;
    const uint32_t* xsrc = ph.pro == PRO_ATTN ? P.att_ll + (size_t)rep * P.att_stride : ph.pro == PRO_NORM ? P.h_ll + (size_t)rep * P.h_stride : P.act_ll + (size_t)rep * P.act_stride;
    // fetch (poll) one run of 16 activations
    auto fetch = [&](int u, float* v) {
        if (ph.pro == PRO_PLAIN) load_run_f32(ph.src, u, v);
        else ll16_poll_run(xsrc, u, tin16, v);
    };
    auto fetch2 = [&](int ua, int ub, float* va, float* vb) {
        if (ph.pro == PRO_PLAIN) { load_run_f32(ph.src, ua, va); load_run_f32(ph.src, ub, vb); }
        else ll16_poll_run2(xsrc, ua, ub, tin16, va, vb);
    };
    auto quantise_store = [&](int u, float* v) {
        float amax = 0.f;
#pragma unroll
        for (int i = 0; i < 16; i++) amax = fmaxf(amax, fabsf(v[i]));
        amax = half16_max(amax);
        // scale rule shared with oracle/oracle.c quantise_act_256: d = amax/127, q = rint(x * (127/amax)); the codes are the
        // same for both layout families, only their place in shared memory differs
        const float d = amax / 127.0f, id = amax > 0.f ? 127.0f / amax : 0.f;
        QRun q; quantise_run(v, id, q);
        if (fam_std) store_run(sm, T_Q4_K, u, q, d);
        if (fam_q6) store_run(s2, T_Q6_K, u, q, d);
    };
    float in_scale = 1.f;                           // factor applied to this phase's dot products in the epilogue
    if (ph.pro == PRO_NORM) {
        // RMSNorm in one pass: a run is multiplied by its norm weights and quantised as soon as it has arrived (int8
        // block quantisation does not depend on a common positive factor); the 1/rms factor is reduced behind the
        // quantisation and applied to the dot products in the epilogue (in_scale).  The norm weights are requested
        // BEFORE the poll (they do not depend on the exchange).
        float v[16], w[16]; float ss = 0.f;
        const bool have = tid < n_runs;
        if (have) {
            const float4* wv = (const float4*)(ph.norm_w + 16 * tid);
#pragma unroll
            for (int i = 0; i < 4; i++) { float4 t = __ldg(wv + i); w[4 * i] = t.x; w[4 * i + 1] = t.y; w[4 * i + 2] = t.z; w[4 * i + 3] = t.w; }
            fetch(tid, v);
#pragma unroll
            for (int i = 0; i < 16; i++) { ss += v[i] * v[i]; v[i] *= w[i]; }
        }
        if (tid == 0) sm.ctrl[0] = 0x7fffffff;                       // inputs have arrived: the producer may run ahead again
        if (tr && tid == 0) tr[1] = tr[2] = gtime();
        WT(1);
        consumer_sync();                                    // red[] and xq free (previous phase's readers done)
        WT(2);
        if (tid < n_runs_pad) {
            if (!have) {
#pragma unroll
                for (int i = 0; i < 16; i++) v[i] = 0.f;
            }
            quantise_store(tid, v);
        }
        for (int u = tid + CONSUMER_THREADS; u < n_runs_pad; u += CONSUMER_THREADS) {
            float t[16];
            if (u < n_runs) {
                fetch(u, t);
                const float4* wv = (const float4*)(ph.norm_w + 16 * u);
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    float4 q4 = __ldg(wv + i);
                    ss += t[4 * i] * t[4 * i] + t[4 * i + 1] * t[4 * i + 1] + t[4 * i + 2] * t[4 * i + 2] + t[4 * i + 3] * t[4 * i + 3];
                    t[4 * i] *= q4.x; t[4 * i + 1] *= q4.y; t[4 * i + 2] *= q4.z; t[4 * i + 3] *= q4.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 16; i++) t[i] = 0.f;
            }
            quantise_store(u, t);
        }
        ss = warp_sum(ss);
        if (lane == 0) sm.red[warp] = ss;
    } else {
        // single pass; a thread's first two runs are polled together (both loads in flight) before the barrier, so the
        // exchange wait overlaps the other warps' tails
        float t[16], t2[16]; const int u2 = tid + CONSUMER_THREADS;
        const bool first = tid < n_runs, second = u2 < n_runs;
        if (second) fetch2(tid, u2, t, t2); else if (first) fetch(tid, t);
        if (tid == 0) sm.ctrl[0] = 0x7fffffff;
        if (tr && tid == 0) tr[1] = tr[2] = gtime();
        WT(1);
        consumer_sync();                                    // xq free (previous phase's readers done)
        WT(2);
        if (tid < n_runs_pad) {
            if (!first) {
#pragma unroll
                for (int i = 0; i < 16; i++) t[i] = 0.f;
            }
            quantise_store(tid, t);
        }
        if (u2 < n_runs_pad) {
            if (!second) {
#pragma unroll
                for (int i = 0; i < 16; i++) t2[i] = 0.f;
            }
            quantise_store(u2, t2);
        }
        for (int u = tid + 2 * CONSUMER_THREADS; u < n_runs_pad; u += CONSUMER_THREADS) {
            if (u < n_runs) fetch(u, t);
            else {
#pragma unroll
                for (int i = 0; i < 16; i++) t[i] = 0.f;
            }
            quantise_store(u, t);
        }
    }
    WT(3);
    consumer_sync();
    WT(4);
    if (ph.pro == PRO_NORM) {
        // all loads first (the loop below over shared memory compiled to 14 dependent load-add pairs), same summation order
        float rr[(CONSUMER_WARPS + 3) / 4 * 4];
#pragma unroll
        for (int i = 0; i < (CONSUMER_WARPS + 3) / 4; i++) { const float4 t = *(const float4*)(sm.red + 4 * i); rr[4 * i] = t.x; rr[4 * i + 1] = t.y; rr[4 * i + 2] = t.z; rr[4 * i + 3] = t.w; }
        float tot = 0.f;
#pragma unroll
        for (int i = 0; i < CONSUMER_WARPS; i++) tot += rr[i];
        in_scale = rsqrtf(tot / (float)ph.K + P.rms_eps);
    }
    if (tr && tid == 0) tr[3] = gtime();
#ifdef B200_TRACE_WAIT
    if (tr && tid == 0) { tr[1] = 0; tr[2] = 0; }
#endif

    // ---- stream the CTA's slice
    int lo, hi; cta_slice(ph.total_pairs, c, G, lo, hi);
    int off = 0;
    for (int o = 0; o < ph.n_ops; o++) {
        const OpDesc& op = ph.op[o];
        const int a = max(lo, off) - off, b = min(hi, off + op.n_pairs) - off; off += op.n_pairs;
        if (a >= b) continue;
        const Smem& sx = (op.type == T_Q6_K) ? s2 : sm;
        // phases with one stage per row block and several row blocks per warp keep the activation slice in registers
        const bool xhold = stage_groups(op.type) == 2 && op.n_grp == 2 && (b - a) > WARP_GROUPS * PAIRS_PER_STAGE;
#define B200_STREAM(T_) do { if (xhold) stream_op<T_, true, TP, TR>(P, op, sm, sx, stage_seq, a, b, pos, warp, lane, seq, in_scale, best, logits_dst, tr); \
                             else stream_op<T_, false, TP, TR>(P, op, sm, sx, stage_seq, a, b, pos, warp, lane, seq, in_scale, best, logits_dst, tr); } while (0)
        constexpr bool TP = (CFG & DK_TP) != 0, TR = (CFG & DK_TRACE) != 0;
        // only the weight types the kernel variant was built for (the host picks a variant that covers the file's types)
        if constexpr ((CFG & DK_Q4K) != 0) { if (op.type == T_Q4_K) B200_STREAM(T_Q4_K); }
        if constexpr ((CFG & DK_Q6K) != 0) { if (op.type == T_Q6_K) B200_STREAM(T_Q6_K); }
        if constexpr ((CFG & DK_Q80) != 0) { if (op.type == T_Q8_0) stream_op<T_Q8_0, false, TP, TR>(P, op, sm, sx, stage_seq, a, b, pos, warp, lane, seq, in_scale, best, logits_dst, tr); }
        if constexpr ((CFG & DK_Q5K) != 0) { if (op.type == T_Q5_K) B200_STREAM(T_Q5_K); }
#undef B200_STREAM
        stage_seq += (uint32_t)(((b - a + PAIRS_PER_STAGE - 1) / PAIRS_PER_STAGE) * stages_per_rb(op.type, op.n_grp));
    }
}

// ------------------------------------------------------------------------------------ the kernel
struct DecodeRun {
    const int* forced; int n_forced;     // tokens fed regardless of argmax (prompt); forced[0] is the first input
    float* logits_all;                   // optional [n_steps][n_vocab]
};

// CFG (DK_* bits): what this variant of the kernel contains.  The host launches the smallest variant that covers the model
// (head size, weight types, tensor parallelism): code the model never runs still costs instruction-cache misses and
// registers at every phase change of a 22 k-instruction kernel (measured: +2.6 % tokens/s for the 8B Q4_K_M file).
template <int CFG>
__global__ void __launch_bounds__(CTA_THREADS, 1)
decode_kernel_t(const DecodeParams P, const DecodeRun R, const uint32_t ring_bytes, const int max_grp) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const Smem sm = carve_smem(smem_raw, ring_bytes, max_grp, P.hown_cap);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, c = blockIdx.x, G = gridDim.x;
    if (tid == 0) {
        for (int i = 0; i < NSLOT; i++) { mbar_init(&sm.full[i], 1); mbar_init(&sm.empty[i], GROUP_WARPS); }
        fence_barrier_init();
        sm.ctrl[0] = 0x7fffffff;
    }
    for (int i = tid; i < P.n_btab; i += CTA_THREADS) sm.btab[i] = P.kv.block_table[i];
    constexpr int PW = (int)(sizeof(Phase) / 4);                        // words per phase descriptor
    {   // schedule -> shared memory: all of a short schedule, or embed / lm_head / the first two layers of a model's
        const uint32_t* src = (const uint32_t*)P.phases; uint32_t* dst = (uint32_t*)sm.phases;
        if (P.sched_layers == 0) { for (int i = tid; i < P.n_phases * PW; i += CTA_THREADS) dst[i] = __ldg(src + i); }
        else {
            for (int i = tid; i < PW; i += CTA_THREADS) { dst[i] = __ldg(src + i); dst[PW + i] = __ldg(src + (size_t)(P.n_phases - 1) * PW + i); }
            for (int i = tid; i < 2 * LAYER_PHASES * PW; i += CTA_THREADS) {
                const int half = i / (LAYER_PHASES * PW), w = i - half * LAYER_PHASES * PW, layer = half % P.sched_layers;
                dst[2 * PW + i] = __ldg(src + (size_t)(1 + layer * LAYER_PHASES) * PW + w);
            }
        }
    }
    build_segs(sm, P, c, G, tid, CTA_THREADS, sm.ring);                 // ends with __syncthreads

    if (warp == CONSUMER_WARPS) {           // ===== producer warp =====
        Producer pr; pr.init(P.max_inflight);
        pr.run(sm, P, sm.ctrl[1], R.n_forced, R.logits_all != nullptr, lane);
        pr.drain(sm);
        return;
    }

    // ===== consumer warps =====
    uint32_t seq = *(volatile unsigned*)P.epoch;          // generation counter: one per executed phase
    const int pos0 = *(volatile int*)P.pos;
    int token = R.n_forced > 0 ? R.forced[0] : *(volatile int*)P.token;
    uint32_t sched_pf = 0;                                  // one word of the next layer's descriptors in flight (see the phase loop)
    uint32_t stage_seq = 0;                                 // stages of this CTA consumed so far (the producer's cur.seq, on the consumer side)
    int hlo, hhi; cta_slice(P.n_embd / 2, c, G, hlo, hhi);  // hidden-state row pairs this CTA owns (== its wo/down slices)
    for (int step = 0; step < P.n_steps; step++) {
        const int pos = pos0 + step;
        const bool need_lm = (step + 1 >= R.n_forced) || R.logits_all;
        float* logits_dst = R.logits_all ? R.logits_all + (size_t)step * P.n_vocab_l : P.logits;
        ArgBest best; best.v = -INFINITY; best.i = 0x7fffffff;
        for (int i = 0; i < P.n_phases; i++) {
            const Phase& ph = sm.phases[phase_slot(P, step, i)];
            if (ph.kind == PH_GEMV && ph.op[0].epi == EPI_LOGITS && !need_lm) continue;
            if (P.sched_layers > 0 && i > 0 && i < P.n_phases - 1 && tid < LAYER_PHASES * PW) {
                // the layer after this one: its descriptors are requested at the layer's first phase and stored at its second
                // (a full phase for the load to land), into the half the previous layer has left (every warp passed this
                // layer's first prologue barrier since)
                const int l = (i - 1) / LAYER_PHASES, j = (i - 1) - l * LAYER_PHASES;
                if (j == 0) sched_pf = __ldg((const uint32_t*)P.phases + (size_t)(1 + ((l + 1) % P.sched_layers) * LAYER_PHASES) * PW + tid);
                else if (j == 1) ((uint32_t*)sm.phases)[(2 + ((step * P.sched_layers + l + 1) & 1) * LAYER_PHASES) * PW + tid] = sched_pf;
            }
            seq++;
            long long* tr = nullptr;
            if constexpr ((CFG & DK_TRACE) != 0) { if (P.trace) tr = P.trace + (((size_t)step * P.n_phases + i) * G + c) * 8; }
            if (tr && tid == 0) { const long long t0 = gtime(); for (int k = 0; k < 7; k++) tr[k] = t0; tr[7] = 0; }
#ifdef B200_WTRACE
            if (lane == 0) s_wt[warp] = (g_wtrace && c == g_wtrace_cta) ? g_wtrace + (((size_t)step * P.n_phases + i) * 16 + warp) * 8 : nullptr;
            __syncwarp();
            WT(0);
#endif
            // this thread is about to poll an exchange: until its inputs have arrived the producer issues nothing beyond
            // the phases the CTA has already entered (slower warps may still be consuming those)
            if (tid == 0 && ph.kind != PH_EMBED) sm.ctrl[0] = step * P.n_phases + i - 1;
#ifdef B200_TRACE_WAIT
            const long long cphase0 = clock64();
#endif
            if (ph.kind == PH_EMBED) {
                // every warp must be done with the previous token (its down-proj epilogue updates hown and publishes
                // this CTA's rows); without this a late RESID publish could overwrite the new token's tagged rows
                consumer_sync();
                if (tid < P.n_rot / 2 && tid < 128) { float sn, cs; sincosf((float)pos * __ldg(P.inv_freq + tid), &sn, &cs); sm.rope[tid] = make_float2(cs, sn); }
                const int ng = n_groups(P.n_embd); const uint32_t t16 = tag16_of(seq);
                for (int e = 2 * hlo + tid; e < 2 * hhi; e += CONSUMER_THREADS) {
                    float v = embed_elem(P.embd, P.embd_type, ng, token, e);
                    if (P.emb_scale != 0.f) v *= P.emb_scale;
                    sm.hown[e - 2 * hlo] = v;
                    st_rep_u32(P.h_ll, P.h_stride, e, ll16_pack(v, t16));
                }
                consumer_sync();
            } else if (ph.kind == PH_ATTN) {
                for (int it = c; it < P.n_head * P.n_splits; it += G) {
                    if constexpr ((CFG & DK_HD128) != 0) { if (P.head_dim == 128) attn_item<4, (CFG & DK_TRACE) != 0>(P, sm, ph.layer, it / P.n_splits, it % P.n_splits, pos, seq - 1, seq, warp, lane, tr); }
                    if constexpr ((CFG & DK_HD64) != 0) { if (P.head_dim != 128) attn_item<2, (CFG & DK_TRACE) != 0>(P, sm, ph.layer, it / P.n_splits, it % P.n_splits, pos, seq - 1, seq, warp, lane, tr); }
                }
                if (tid == 0) sm.ctrl[0] = 0x7fffffff;                 // CTAs without an attention item polled nothing
            } else {
                consume_phase<CFG>(P, ph, sm, stage_seq, c, G, pos, tid, seq, best, logits_dst, tr);
                if (ph.op[0].epi == EPI_LOGITS) {          // publish this CTA's argmax candidate
                    consumer_sync();
                    if (lane == 0) { sm.red[2 * warp] = best.v; sm.red[2 * warp + 1] = __int_as_float(best.i); }
                    consumer_sync();
                    if (tid == 0) {
                        float bv = -INFINITY; int bi = 0x7fffffff;
                        for (int w = 0; w < CONSUMER_WARPS; w++) {
                            const float vv = sm.red[2 * w]; const int ix = __float_as_int(sm.red[2 * w + 1]);
                            if (vv > bv || (vv == bv && ix < bi)) { bv = vv; bi = ix; }
                        }
                        st_relaxed_v2(P.cand_ll + 2 * c, __float_as_uint(bv), seq); st_relaxed_v2(P.cand_ll + 2 * c + 1, (uint32_t)bi, seq);
                        if constexpr ((CFG & DK_TP) != 0) for (int r = 0; r < P.tp_world; r++) if (r != P.tp_rank) {      // vocabulary shards: every rank sees every rank's candidates
                            uint2* dst = P.cand_peer[r] + (size_t)P.tp_rank * 2 * G + 2 * c;
                            st_sys_v2(dst, __float_as_uint(bv), seq); st_sys_v2(dst + 1, (uint32_t)bi, seq);
                        }
                    }
                }
            }
#ifdef B200_TRACE_WAIT
            if (tr && tid == 0) { tr[6] = gtime(); tr[4] = tr[0] + (clock64() - cphase0); }   // tr[4]-tr[0] = SM cycles of the phase; tr[7] = warp 0's wait-for-data ns
#else
            if (tr && tid == 0) tr[6] = gtime();
#endif
            WT(7);
        }
        // ---- next token: forced, or argmax over the per-CTA candidates (every CTA redundantly)
        int next;
        if (step + 1 < R.n_forced) next = R.forced[step + 1];
        else {
            float bv = -INFINITY; int bi = 0x7fffffff;
            for (int k = tid; k < G; k += CONSUMER_THREADS) {
                float t[2]; ll32_poll<2>(P.cand_ll + 2 * k, seq, t);
                const float vv = t[0]; const int ix = __float_as_int(t[1]);
                if (vv > bv || (vv == bv && ix < bi)) { bv = vv; bi = ix; }
            }
            if constexpr ((CFG & DK_TP) != 0) if (P.tp_world > 1) for (int k = tid; k < P.tp_world * G; k += CONSUMER_THREADS) {
                if (k / G == P.tp_rank) continue;                       // own candidates were read above
                const uint2* src = P.cand_peer[P.tp_rank] + 2 * (size_t)k;
                uint2 w0, w1; unsigned spins = *(volatile unsigned*)P.tp_error ? TP_SPIN_BUDGET : 0u;
                for (;;) {
                    w0 = ld_sys_v2(src); w1 = ld_sys_v2(src + 1);
                    if (w0.y == seq && w1.y == seq) break;
                    if (++spins > TP_SPIN_BUDGET) { *P.tp_error = 1u; w0.x = __float_as_uint(-INFINITY); w1.x = 0x7fffffffu; break; }
                }
                const float vv = __uint_as_float(w0.x); const int ix = (int)w1.x;
                if (vv > bv || (vv == bv && ix < bi)) { bv = vv; bi = ix; }
            }
            // Lanes leave the polls above at different times and most warps skip them.  Without this barrier the variants of the
            // kernel that have no tensor-parallel block between the poll and the shuffles reduced stale registers (found by
            // test_greedy_stream_matches_oracle when the variants were introduced).
            __syncwarp();
#pragma unroll
            for (int o = 16; o; o >>= 1) {
                const float vv = __shfl_xor_sync(0xffffffffu, bv, o); const int ix = __shfl_xor_sync(0xffffffffu, bi, o);
                if (vv > bv || (vv == bv && ix < bi)) { bv = vv; bi = ix; }
            }
            consumer_sync();
            if (lane == 0) { sm.red[2 * warp] = bv; sm.red[2 * warp + 1] = __int_as_float(bi); }
            consumer_sync();
            bv = -INFINITY; bi = 0x7fffffff;
            for (int w = 0; w < CONSUMER_WARPS; w++) {
                const float vv = sm.red[2 * w]; const int ix = __float_as_int(sm.red[2 * w + 1]);
                if (vv > bv || (vv == bv && ix < bi)) { bv = vv; bi = ix; }
            }
            next = bi < P.n_vocab ? bi : 0;                    // all-NaN logits: stay inside the vocabulary (the token indexes the embedding table)
        }
        if (c == 0 && tid == 0) P.out_tokens[step] = next;
        token = next;
        __threadfence();     // once per token: orders this CTA's KV-page stores before anything a later step publishes
    }
    if (c == 0 && tid == 0) { *P.pos = pos0 + P.n_steps; *P.token = token; *P.epoch = seq; }
}

// ------------------------------------------------------------------------------------ repack (K9) and test kernels
// One thread per 16 output bytes; source bytes gathered through src_byte().  parity >= 0 interleaves:
// raw row r goes to engine row 2r + parity (gate/up fusion).  Rows >= n_src are zero (odd-N padding).
// src_stride: bytes between source rows (0 = dense rows of K elements); a K-slice of a wider matrix passes the full row size.
__global__ void repack_kernel(const uint8_t* __restrict__ raw, uint8_t* __restrict__ out, int type, long long N, long long n_src, int K, int parity, size_t src_stride) {
    const int gb = group_bytes(type), ng = n_groups(K); const size_t rrb = src_stride ? src_stride : raw_row_bytes(type, K);
    const long long chunks_per_row = (long long)ng * gb / 16, total = N * chunks_per_row;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long row = i / chunks_per_row; const int rem = (int)(i % chunks_per_row);
        const int g = rem / (gb / 16), o = (rem % (gb / 16)) * 16;
        uint8_t b[16];
#pragma unroll
        for (int k = 0; k < 16; k++) { long long s = row < n_src ? src_byte(type, K, g, o + k) : -1; b[k] = s < 0 ? 0 : raw[(size_t)row * rrb + s]; }
        if ((type == T_Q4_K || type == T_Q5_K) && o % block_bytes(type) == 0) { uint8_t e[12]; k4_reencode(b + 4, e); for (int k = 0; k < 12; k++) b[4 + k] = e[k]; }
        const long long orow = parity < 0 ? row : 2 * row + parity;
        uint8_t* dst = out + ((((size_t)(orow >> 1)) * ng + g) * 2 + (orow & 1)) * gb + o;
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
