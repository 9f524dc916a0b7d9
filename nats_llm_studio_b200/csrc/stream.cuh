// stream.cuh — the weight-streaming core of the batch-1 decode path (kernels K1, K3, K4, K6, K11 of
// SURVEY.md §2.3 fused into one persistent kernel; replaces the arithmetic behind
// /root/reference/nats_llm_studio.go:158-179, which the reference delegates to LM Studio).
//
// Design (B200-first, HBM-bound work):
//   * one persistent cooperative grid, one CTA per SM, 17 warps: warp 16 is the PRODUCER, warps 0..15
//     are CONSUMERS in two groups of 8 that take alternate row blocks (two stages consumed at once).
//   * the producer walks the token's static weight schedule and pulls each CTA's slice of every
//     matrix through a shared-memory byte ring with 1-D bulk-async copies (cp.async.bulk, the TMA
//     engine; SASS UBLKCP) completing on mbarriers.  It never waits on data dependencies, only on
//     ring space, so HBM keeps streaming across phase boundaries, grid barriers and prologues.
//   * consumers quantise the activation vector once per phase into shared memory (int8 per
//     256-block, the way ggml quantises activations before a quantised mat-vec), then each warp
//     owns one ROW PAIR of a stage: every lane reads its 64 weights per row with conflict-free
//     128-bit shared loads, unpacks nibbles with LOP3 and accumulates with dp4a (IDP.4A), integer
//     per super-block, fp32 across super-blocks.
//   * epilogues are fused: RoPE + KV-page write, residual add, SiLU*mul, logit scale + argmax.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>
#include "quant.cuh"

namespace b200 {

// decode_kernel_t<CFG> variants (decode.cuh): what a variant contains
constexpr int DK_HD128 = 1, DK_HD64 = 2, DK_TP = 4, DK_Q4K = 8, DK_Q5K = 16, DK_Q6K = 32, DK_Q80 = 64, DK_TRACE = 128;   // DK_TRACE: B200_TRACE stamps
constexpr int DK_TYPES = DK_Q4K | DK_Q5K | DK_Q6K | DK_Q80, DK_ALL = DK_HD128 | DK_HD64 | DK_TP | DK_TYPES | DK_TRACE;

#ifndef B200_GROUP_WARPS
#define B200_GROUP_WARPS 7      // 14 consumer warps + the producer = 15 warps: at most 4 warps per scheduler, so 128 registers per thread
                                // (16 + 1 warps cap the kernel at 96 registers: measured 1.97 ms/token on the 8B model against 1.89)
#endif
constexpr int GROUP_WARPS = B200_GROUP_WARPS;   // warps that share one stage (one row pair each)
#ifndef B200_WARP_GROUPS
#define B200_WARP_GROUPS 2
#endif
constexpr int WARP_GROUPS = B200_WARP_GROUPS;           // groups take alternate row blocks, so two stages are consumed concurrently
constexpr int CONSUMER_WARPS = GROUP_WARPS * WARP_GROUPS;
constexpr int CONSUMER_THREADS = CONSUMER_WARPS * 32;
constexpr int CTA_THREADS = CONSUMER_THREADS + 32;
constexpr int PAIRS_PER_STAGE = GROUP_WARPS;
constexpr int NSLOT = 16;                // mbarrier pairs in flight
#ifndef B200_XREP
#define B200_XREP 4
#endif
constexpr int XREP = B200_XREP;          // copies of every broadcast exchange vector (an L2 slice serves ~1 sector/clk: one copy read by
                                         // every CTA costs ~0.6 us per line; CTA c reads copy c % XREP).  Measured 2, 3, 4: 611 tok/s, 8: 608, 16: 576
// groups (of 2048 elements) per stage: one bulk copy moves a pair's run of `stage_groups` consecutive groups, which
// are contiguous in HBM — fewer, larger copies (the single producer warp issues them one by one)
#ifndef B200_STAGE_GROUPS
#define B200_STAGE_GROUPS 2
#endif
__host__ __device__ constexpr int stage_groups(int type) { return type == 8 /*Q8_0*/ ? 1 : B200_STAGE_GROUPS; }
constexpr int MAX_STAGE_BYTES = PAIRS_PER_STAGE * 2 * 2 * 1680;   // Q6_K, two groups

// ------------------------------------------------------------------------------------ PTX
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(CONSUMER_THREADS) : "memory"); }
__device__ __forceinline__ int dp4a_ss(int a, int b, int c) { return __dp4a(a, b, c); }
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {   // unsigned bytes x signed bytes
    int d; asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ------------------------------------------------------------------------------------ descriptors
enum Epilogue : int { EPI_STORE = 0, EPI_ROPE_Q, EPI_ROPE_K, EPI_V, EPI_RESID, EPI_SWIGLU, EPI_LOGITS };
enum Prologue : int { PRO_PLAIN = 0, PRO_NORM, PRO_ATTN, PRO_ACT };   // PLAIN: fp32 src (GEMV entry point); others poll tagged vectors
enum PhaseKind : int { PH_GEMV = 0, PH_ATTN, PH_EMBED };

struct OpDesc {                 // one streamed matrix (engine layout, quant.cuh)
    const uint8_t* w;
    int type, n_pairs, n_grp, epi;
    int n_rows;                 // valid rows (2*n_pairs, or one less when an odd matrix was padded)
    float* out;                 // EPI_STORE destination
    float scale;                // output scale (residual_scale, 1/logit_scale, 1)
    int layer;                  // for KV epilogues
    int tp_half;                // EPI_RESID under tensor parallelism: which half of the peers' receive buffers (0 = wo, 1 = down)
    int pad_;
};

struct Phase {
    int kind, n_ops, total_pairs, K;
    int pro; int layer;
    const float* src;           // PRO_PLAIN source vector — fp32 [K] (kernel-level GEMV entry point)
    const float* norm_w;        // PRO_NORM weight
    OpDesc op[3];
};

struct KvLayout {               // paged KV cache, fp16
    __half* pages;              // [n_layer][n_pages][2][n_head_kv][16][head_dim]: layer-major, so that the pages one layer's
                                // attention touches lie in one compact region (n_pages x 64 KB for 8 kv heads of 128) instead
                                // of being strided by the size of a whole page across all layers (2 MB: a TLB miss per page)
    const int* block_table;     // [max_seqs][pages_per_seq]; the batch-1 decode kernel uses sequence slot 0
    int n_layer, n_head_kv, head_dim, pages_per_seq, n_pages;
    __device__ __forceinline__ __half* at_page(int page, int layer, int kv, int kvh, int pos) const {
        return pages + (((((size_t)layer * n_pages + page) * 2 + kv) * n_head_kv + kvh) * 16 + (pos & 15)) * head_dim;
    }
    __device__ __forceinline__ __half* at_slot(int slot, int layer, int kv, int kvh, int pos) const {
        return at_page(block_table[slot * pages_per_seq + (pos >> 4)], layer, kv, kvh, pos);
    }
    __device__ __forceinline__ __half* at(int layer, int kv, int kvh, int pos) const { return at_page(block_table[pos >> 4], layer, kv, kvh, pos); }
};

struct DecodeParams {
    const Phase* phases; int n_phases;
    int sched_layers;           // > 0: the schedule is [embed][LAYER_PHASES x sched_layers][lm_head]; shared memory then keeps only the
                                //      phases of two layers (the current one and the next, fetched while the current one runs).
                                // 0: any schedule of at most SCHED_SLOTS phases, all resident (kernel-level GEMV entry point)
    int max_segs;               // capacity of the per-CTA slice list (number of (phase, matrix) pairs of a token)
    // model scalars
    int n_embd, n_head, n_head_kv, head_dim, n_rot, n_vocab, n_splits;
    int n_vocab_l;              // rows of this rank's lm_head (== n_vocab on one GPU): row stride of logits_all
    float rms_eps, kq_scale, emb_scale;
    const float* inv_freq;      // [n_rot/2] = base^(-2i/n_rot) / freq_factor_i
    // embedding
    const uint8_t* embd; int embd_type;   // engine layout
    // buffers
    float* logits;
    // tagged exchange vectors (decode.cuh): LL16 = uint32 {tag16, fp16}, LL32 = uint2 {bits, tag32}
    uint32_t* h_ll; uint32_t* att_ll; uint32_t* act_ll; uint32_t* kv_ll;   // [XREP][n_embd], [XREP][n_head*head_dim], [XREP][n_ff], [2*n_head_kv*head_dim]
    int h_stride, att_stride, act_stride;                                   // words between copies
    uint2* q_ll; uint2* part_ll; uint2* cand_ll;                            // [n_head*head_dim], [n_head][16][head_dim+2], [2*grid]
    int hown_cap;                                                           // floats of owned hidden-state rows per CTA
    int n_btab;                                                             // block-table entries copied to shared memory
    KvLayout kv;
    // run state (device memory)
    int* pos; int* token; int* out_tokens; unsigned* bar_ctr; unsigned* epoch;
    int n_steps; int want_argmax; int max_inflight;
    int quiet;                  // 1: the producer holds back copies of later phases while the CTA polls an exchange (see Smem::ctrl)
    // tensor parallelism inside the kernel (K10): this rank's shard runs the same schedule; the partial sums of the two
    // row-parallel products of a layer and the argmax candidates are exchanged as tagged words over NVLink peer memory
    int tp_rank, tp_world;      // tp_world == 1: single GPU (nothing below is used)
    uint2* rx_peer[8];          // rank r's receive buffer as mapped in THIS process: [2 (wo | down)][tp_world (source rank)][n_embd] LL32 words
    uint2* cand_peer[8];        // rank r's candidate buffer: [tp_world (source rank)][2 * grid] LL32 words
    int vocab_base;             // first vocabulary row of this rank's lm_head slice
    unsigned* tp_error;         // set when a peer did not answer within the spin budget (the launch then ends with garbage instead of hanging)
    long long* trace;           // optional [n_steps][n_phases][grid][4]: clock64 at phase start / prologue end / phase end, smid (debug)
};

// ------------------------------------------------------------------------------------ smem carve-up
struct Smem {
    uint8_t* ring; uint32_t ring_bytes;
    int8_t* xq; int16_t* xbs; float* xd;     // activation vector, engine-permuted
    uint64_t* full; uint64_t* empty;
    float* red;                               // 64 floats of reduction scratch
    uint32_t* q_end;                          // producer's in-flight stage sizes [NSLOT]
    uint32_t* q_off;                          // ring offset of the stage in each barrier slot [NSLOT] (written by the producer before it arms the slot)
    float* hown;                              // fp32 hidden-state rows this CTA owns (residual stream never leaves the SM)
    float2* rope;                             // (cos, sin) of this token's RoPE angles [128]
    int* btab;                                // the sequence's KV block table (page ids)
    Phase* phases;                            // SCHED_SLOTS phase descriptors (descriptor reads are on every phase's critical path; an
                                              // L2 round trip each would cost microseconds): [0] embed, [1] lm_head, [2..] two layers
    struct Seg* segs;                         // this CTA's slices of the token's streamed matrices, in schedule order (producer)
    volatile int* ctrl;                       // [0] quiet_id: phase id (step * n_phases + phase) up to which copies may be issued
                                              //     while the consumers poll an exchange; INT_MAX = no restriction.  [1] slices in segs[]
};
constexpr int LAYER_PHASES = 5;               // QKV, attention, wo, gate/up, down
constexpr int SCHED_SLOTS = 2 + 2 * LAYER_PHASES;
// shared-memory slot of phase i of token `step`: layers alternate between the two halves by their running number
__device__ __forceinline__ int phase_slot(const DecodeParams& P, int step, int i) {
    if (P.sched_layers == 0) return i;
    if (i == 0) return 0;
    if (i == P.n_phases - 1) return 1;
    const int l = (i - 1) / LAYER_PHASES, j = (i - 1) - l * LAYER_PHASES;
    return 2 + ((step * P.sched_layers + l) & 1) * LAYER_PHASES + j;
}

// One streamed slice: pairs [0, n_pairs) starting at src, n_grp groups per row, of phase `phase` of the token.
struct Seg { const uint8_t* src; unsigned short n_pairs; unsigned char n_grp, type; unsigned short phase; unsigned char is_lm, pad; };
static_assert(sizeof(Seg) == 16, "Seg layout");
constexpr int BTAB_MAX = 1024;           // KV pages per sequence cached in shared memory (16384 positions)
constexpr int XG_MAX = 16;               // activation groups (2048 elements each) the fixed part of the layout has scale / sum slots for
// Shared-memory layout.  Everything the stream loop and the phase changes touch sits at a COMPILE-TIME offset from the start
// of the dynamic array (barriers, ring bookkeeping, schedule, block table, activation scales and sums, then the ring itself):
// with the ring first and the rest behind it every one of these addresses depended on two launch parameters and was
// re-derived (LDC, shifts, adds) inside the stage loop.  Only the int8 activation vector, the owned hidden rows and the
// slice list, whose sizes depend on the model, follow the ring.
struct SmemLayout {
    static constexpr uint32_t FULL = 0, EMPTY = FULL + NSLOT * 8, Q_END = EMPTY + NSLOT * 8, Q_OFF = Q_END + NSLOT * 4, CTRL = Q_OFF + NSLOT * 4,
                              RED = CTRL + 64, ROPE = RED + 64 * 4, BTAB = ROPE + 128 * 8, PHASES = BTAB + BTAB_MAX * 4,
                              XD = (PHASES + SCHED_SLOTS * (uint32_t)sizeof(Phase) + 15) / 16 * 16, XBS = XD + XG_MAX * 32, RING = (XBS + XG_MAX * 256 + 127) / 128 * 128;
};
__host__ __device__ inline size_t smem_tail_bytes(int max_grp, int hown_cap, int max_segs) {
    return (size_t)max_grp * 2048 + ((size_t)hown_cap * 4 + 15) / 16 * 16 + (size_t)max_segs * sizeof(Seg) + 64;
}
__device__ inline Smem carve_smem(uint8_t* base, uint32_t ring_bytes, int max_grp, int hown_cap) {   // hown_cap: multiple of 4
    using L = SmemLayout;
    Smem s;
    s.full = (uint64_t*)(base + L::FULL); s.empty = (uint64_t*)(base + L::EMPTY); s.q_end = (uint32_t*)(base + L::Q_END); s.q_off = (uint32_t*)(base + L::Q_OFF);
    s.ctrl = (volatile int*)(base + L::CTRL); s.red = (float*)(base + L::RED); s.rope = (float2*)(base + L::ROPE); s.btab = (int*)(base + L::BTAB);
    s.phases = (Phase*)(base + L::PHASES); s.xd = (float*)(base + L::XD); s.xbs = (int16_t*)(base + L::XBS);
    s.ring = base + L::RING; s.ring_bytes = ring_bytes;
    uint8_t* p = s.ring + ring_bytes;
    s.xq = (int8_t*)p; p += (size_t)max_grp * 2048;
    s.hown = (float*)p; p += ((size_t)hown_cap * 4 + 15) / 16 * 16;
    s.segs = (Seg*)p;
    return s;
}

// Deterministic ring allocator state; producer and consumers evolve identical copies.
struct RingCursor {
    uint32_t seq;        // stage sequence number
    uint32_t off;        // byte offset where the next stage would start
    __device__ __forceinline__ void init() { seq = 0; off = 0; }
    // stages never straddle the wrap point: a stage that does not fit behind `off` starts at 0 and leaves a gap
    __device__ __forceinline__ uint32_t place(uint32_t bytes, uint32_t ring_bytes, uint32_t& gap) {
        gap = 0;
        if (off + bytes > ring_bytes) { gap = ring_bytes - off; off = 0; }
        const uint32_t at = off; off += bytes; return at;
    }
};

// slice of the phase's concatenated pair space owned by CTA c of G
__device__ __forceinline__ void cta_slice(int total, int c, int G, int& lo, int& hi) {   // total * G < 2^32 (pairs x SMs)
    lo = (int)((uint32_t)total * (uint32_t)c / (uint32_t)G); hi = (int)((uint32_t)total * (uint32_t)(c + 1) / (uint32_t)G);
}

// ------------------------------------------------------------------------------------ producer
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}

// Stage order inside one slice (n_rb row blocks of 8 pairs, spr stages per row block): row blocks are taken two at a
// time — one for each consumer warp group — and their stages alternate, so that both groups always have a stage in the
// ring (with all of block A's stages ahead of block B's, group B would sit idle until A is nearly done: the ring holds
// about four stages).  With one stage per row block this is simply ascending row-block order.
__host__ __device__ inline int stages_per_rb(int type, int n_grp) { const int sg = stage_groups(type); return (n_grp + sg - 1) / sg; }
__host__ __device__ inline int seg_stage_index(int rbi, int gi, int n_rb, int spr) {       // position of (row block rbi, stage gi) in the slice's order
    const int pp = rbi >> 1, nb = min(2, n_rb - 2 * pp);
    return pp * 2 * spr + gi * nb + (rbi & 1);
}
static_assert(WARP_GROUPS == 2, "the stage order pairs row blocks for two consumer warp groups");

// Position in the CTA's stage sequence: (token step, slice, row-block pair, stage, which block of the pair).  The ring cursor
// and the L2-prefetch cursor are two of these walking the same list.
struct SegIter {
    int step, si, pp, gi, w; uint32_t n;              // n: stages passed so far
    __device__ void init() { step = 0; si = 0; pp = 0; gi = 0; w = 0; n = 0; }
    // current slice, skipping empty ones and lm_head slices of steps that do not need logits; false = past the last step
    __device__ bool get(const Smem& sm, int n_seg, int n_steps, int n_forced, bool logits_all, Seg& s) {
        for (;;) {
            if (step >= n_steps) return false;
            if (si >= n_seg) { step++; si = 0; pp = 0; gi = 0; w = 0; continue; }
            s = sm.segs[si];
            const bool need_lm = (step + 1 >= n_forced) || logits_all;
            if (s.n_pairs == 0 || (s.is_lm && !need_lm)) { si++; pp = 0; gi = 0; w = 0; continue; }
            return true;
        }
    }
    __device__ void advance(const Seg& s) {
        const int n_rb = (s.n_pairs + PAIRS_PER_STAGE - 1) / PAIRS_PER_STAGE, spr = stages_per_rb(s.type, s.n_grp);
        n++;
        if (++w >= min(2, n_rb - 2 * pp)) { w = 0; if (++gi >= spr) { gi = 0; if (2 * ++pp >= n_rb) { pp = 0; si++; } } }
    }
};
struct StageDesc { const uint8_t* src; uint32_t cb; int np; };            // lane's source, bytes per pair, pairs in the stage
__device__ __forceinline__ StageDesc stage_of(const Seg& s, const SegIter& it, int lane) {
    StageDesc d; const uint32_t pu = 2u * group_bytes(s.type); const int sg = stage_groups(s.type);
    const int rb = (2 * it.pp + it.w) * PAIRS_PER_STAGE, g0 = it.gi * sg;
    d.np = min(PAIRS_PER_STAGE, s.n_pairs - rb);
    d.cb = min(sg, s.n_grp - g0) * pu;
    d.src = s.src + ((size_t)(rb + min(lane, d.np - 1)) * s.n_grp + g0) * pu;
    return d;
}

// The CTA's slices of one token, in schedule order: built once per launch by the whole CTA straight from the schedule in
// global memory (phase i by thread i: count, prefix over the counts, write).  `scratch`: n_phases bytes (the ring, still idle).
__device__ inline void build_segs(const Smem& sm, const DecodeParams& P, int c, int G, int tid, int n_threads, unsigned char* scratch) {
    auto slices = [&](int i, Seg* out) {
        const Phase& ph = P.phases[i];
        if (ph.kind != PH_GEMV) return 0;
        int lo, hi; cta_slice(ph.total_pairs, c, G, lo, hi);
        int off = 0, n = 0;
        for (int o = 0; o < ph.n_ops; o++) {
            const OpDesc& op = ph.op[o];
            const int a = max(lo, off) - off, b = min(hi, off + op.n_pairs) - off; off += op.n_pairs;
            if (a >= b) continue;
            if (out) {
                Seg sg; sg.src = op.w + (size_t)a * op.n_grp * 2u * group_bytes(op.type); sg.n_pairs = (unsigned short)(b - a); sg.n_grp = (unsigned char)op.n_grp;
                sg.type = (unsigned char)op.type; sg.is_lm = op.epi == EPI_LOGITS; sg.phase = (unsigned short)i; sg.pad = 0;
                out[n] = sg;
            }
            n++;
        }
        return n;
    };
    for (int i = tid; i < P.n_phases; i += n_threads) scratch[i] = (unsigned char)slices(i, nullptr);
    __syncthreads();
    for (int i = tid; i < P.n_phases; i += n_threads) {
        int off = 0; for (int k = 0; k < i; k++) off += scratch[k];
        if (off + scratch[i] <= P.max_segs) slices(i, sm.segs + off);
        if (i == P.n_phases - 1) sm.ctrl[1] = min(off + scratch[i], P.max_segs);
    }
    __syncthreads();
}

struct Producer {
    RingCursor cur; uint32_t used, oldest;            // bytes held by unreleased stages (incl. gaps); oldest unreleased stage
    uint32_t landed, max_unlanded;                    // stages whose bytes have all arrived; cap on issued-but-unlanded stages
    __device__ void init(int inflight) { cur.init(); used = 0; oldest = 0; landed = 0; max_unlanded = (uint32_t)max(1, inflight); }

    // Whole producer warp.  Walks the stage sequence of all n_steps tokens; never waits on data dependencies, only on ring
    // space — except that, while the consumers poll an exchange (ctrl[0] = id of the last phase they completed), copies of
    // later phases are held back: the poll's replies would queue behind this SM's in-flight bulk data, and what is being
    // fetched at that moment is never the next thing the consumers need.  (Pulling the stages beyond the ring into L2 with
    // cp.async.bulk.prefetch.L2 while blocked on ring space was built and measured: no gain, removed.)
    __device__ void run(const Smem& sm, const DecodeParams& P, int n_seg, int n_forced, bool logits_all, int lane) {
        SegIter it; it.init();
        Seg s;
        while (it.get(sm, n_seg, P.n_steps, n_forced, logits_all, s)) {
            const StageDesc d = stage_of(s, it, lane);
            const uint32_t bytes = d.np * d.cb, slot = cur.seq % NSLOT;
            const int phase_id = it.step * P.n_phases + s.phase;
            // (0) keep the number of un-landed stages small: other L2 accesses of this SM queue behind its outstanding bulk bytes
            while (cur.seq - landed >= max_unlanded) { mbar_wait(&sm.full[landed % NSLOT], (landed / NSLOT) & 1); landed++; }
            // (1) barrier slot reuse and (2) ring space (the stage plus the gap it would leave at the wrap point); (3) exchange quiet
            const uint32_t gp = (cur.off + bytes > sm.ring_bytes) ? sm.ring_bytes - cur.off : 0;
            for (;;) {
                const bool quiet = P.quiet && phase_id > sm.ctrl[0];
                const bool need_release = (cur.seq - oldest >= NSLOT) || (used + gp + bytes > sm.ring_bytes);
                if (!quiet && !need_release) break;
                if (need_release && mbar_test(&sm.empty[oldest % NSLOT], (oldest / NSLOT) & 1)) { used -= sm.q_end[oldest % NSLOT]; oldest++; continue; }
                __nanosleep(64);
            }
            uint32_t gap; const uint32_t soff = cur.place(bytes, sm.ring_bytes, gap);
            used += gap + bytes;
            __syncwarp();                 // every lane is done reading q_end[] of released stages
            if (lane == 0) { sm.q_end[slot] = gap + bytes; sm.q_off[slot] = soff; mbar_expect_tx(&sm.full[slot], bytes); }
            __syncwarp();
            if (lane < d.np) bulk_g2s(sm.ring + soff + lane * d.cb, d.src, d.cb, &sm.full[slot]);
            cur.seq++;
            it.advance(s);
        }
    }
    __device__ void release_oldest(const Smem& sm) {
        const uint32_t slot = oldest % NSLOT;
        mbar_wait(&sm.empty[slot], (oldest / NSLOT) & 1);
        used -= sm.q_end[slot]; oldest++;
    }
    // before the CTA exits every issued copy must have landed and been consumed
    __device__ void drain(const Smem& sm) { while (oldest < cur.seq) release_oldest(sm); }
};

// ------------------------------------------------------------------------------------ activation prologue
// Quantise 16 consecutive activations (one "run") and store them engine-permuted.
// Scale rule (shared with oracle/oracle.c quantise_act_256): d = amax/127, q = rint(x * (127/amax)).
// A quantised run: 16 int8 codes in four words and their sum.  cvt.pack.sat saturates to [-128, 127]; the scale rule maps
// |x| <= amax to |x * id| <= 127 (+ 1 ulp), so -128 is never produced and the result equals the oracle's clamp to +-127.
struct QRun { uint32_t w[4]; int bs; };
__device__ __forceinline__ uint32_t pack_s8x2(int hi, int lo, uint32_t upper) {      // (upper << 16) | (s8(hi) << 8) | s8(lo)
    uint32_t d; asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(hi), "r"(lo), "r"(upper)); return d;
}
__device__ __forceinline__ void quantise_run(const float v[16], float id, QRun& q) {
    q.bs = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const int a0 = __float2int_rn(v[4 * i] * id), a1 = __float2int_rn(v[4 * i + 1] * id), a2 = __float2int_rn(v[4 * i + 2] * id), a3 = __float2int_rn(v[4 * i + 3] * id);
        q.w[i] = pack_s8x2(a1, a0, pack_s8x2(a3, a2, 0u));
        q.bs = __dp4a((int)q.w[i], 0x01010101, q.bs);
    }
}
__device__ __forceinline__ void store_run(const Smem& sm, int type, int u_global, const QRun& q, float d) {
    int grp = u_global / 128, u = u_global % 128, L, plane; act_slot(type, u, L, plane);
    *(uint4*)(sm.xq + (size_t)grp * 2048 + plane * 512 + L * 16) = make_uint4(q.w[0], q.w[1], q.w[2], q.w[3]);
    sm.xbs[grp * 128 + L * 4 + plane] = (int16_t)q.bs;
    if ((u_global & 15) == 0) sm.xd[u_global / 16] = d;
}

__device__ __forceinline__ float half16_max(float a) {   // max over the aligned 16-lane group
#pragma unroll
    for (int o = 8; o; o >>= 1) a = fmaxf(a, __shfl_xor_sync(0xffffffffu, a, o));
    return a;
}

__device__ __forceinline__ void load_run_f32(const float* __restrict__ src, int u, float v[16]) {
    const float4* p = (const float4*)(src + 16 * u);
#pragma unroll
    for (int i = 0; i < 4; i++) { float4 t = __ldcg(p + i); v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w; }
}

// ------------------------------------------------------------------------------------ per-type lane kernels
// Each computes, for lane L of the warp, the contribution of its 64 weights of ONE row's group to
// that row's dot product, already multiplied by the activation block scale dx.
struct XRegs { uint4 p[4]; int bs[4]; float dx; };

__device__ __forceinline__ void load_x(const Smem& sm, int g, int lane, XRegs& x) {
    const int8_t* xb = sm.xq + (size_t)g * 2048 + lane * 16;
#pragma unroll
    for (int i = 0; i < 4; i++) x.p[i] = *(const uint4*)(xb + i * 512);
    uint2 b = *(const uint2*)(sm.xbs + g * 128 + lane * 4);
    x.bs[0] = (int)(int16_t)(b.x & 0xffff); x.bs[1] = (int)(int16_t)(b.x >> 16);
    x.bs[2] = (int)(int16_t)(b.y & 0xffff); x.bs[3] = (int)(int16_t)(b.y >> 16);
    x.dx = sm.xd[g * 8 + lane / 4];
}
__device__ __forceinline__ uint32_t u4_get(const uint4& v, int i) { return i == 0 ? v.x : i == 1 ? v.y : i == 2 ? v.z : v.w; }

// scale / min of sub-blocks 2j, 2j+1 from the engine's per-lane 24-bit fields (quant.cuh k4_reencode)
__device__ __forceinline__ void k4_scales(const uint4& hdr, int j, int& sc0, int& sc1, int& m0, int& m1) {
    const uint32_t lo = j < 2 ? hdr.y : (j == 2 ? hdr.z : hdr.w), hi = j < 2 ? hdr.z : hdr.w;
    const uint32_t f = __funnelshift_r(lo, hi, (32 - 8 * j) & 31);          // bits [24j, 24j+24) of the 96-bit scale field
    sc0 = f & 63; sc1 = (f >> 6) & 63; m0 = (f >> 12) & 63; m1 = (f >> 18) & 63;
}

// The lane's weights of one row's group, in registers: loading (load_w) apart from the arithmetic (dot_w).  (Putting all
// reads of a stage in flight ahead of the arithmetic was tried twice — ordered loads of the whole stage (spills), and two
// WRegs buffers loaded one unit ahead: ptxas sinks every LDS.128 back to just before its first use whatever the source
// order, to stay inside 128 registers — DESIGN.md §3.1.)
struct WRegs { uint4 v[5]; uint32_t s; };
__device__ __forceinline__ uint4 lds128(const void* p) { return *(const uint4*)p; }
__device__ __forceinline__ uint32_t lds32(const void* p) { return *(const uint32_t*)p; }

template <int TYPE>
__device__ __forceinline__ void load_w(const uint8_t* grp, int lane, WRegs& w) {
    if constexpr (TYPE == T_Q4_K) {
        const uint8_t* blk = grp + (lane >> 2) * 144; const int j = lane & 3;
        w.v[0] = lds128(blk); w.v[1] = lds128(blk + 16 + 32 * j); w.v[2] = lds128(blk + 32 + 32 * j);      // d, dmin, scales[12] | 32 + 32 nibbles
    } else if constexpr (TYPE == T_Q5_K) {
        const uint8_t* blk = grp + (lane >> 2) * 176; const int j = lane & 3;
        w.v[0] = lds128(blk); w.v[1] = lds128(blk + 16); w.v[2] = lds128(blk + 32);                         // header | qh[0..31]
        w.v[3] = lds128(blk + 48 + 32 * j); w.v[4] = lds128(blk + 64 + 32 * j);
    } else if constexpr (TYPE == T_Q6_K) {
        w.v[0] = lds128(grp + lane * 16); w.v[1] = lds128(grp + 512 + lane * 16); w.v[2] = lds128(grp + 1024 + lane * 16);
        w.v[3].x = lds32(grp + 1536 + lane * 4);
        w.s = (uint32_t)*(const unsigned short*)(grp + 1664 + (lane >> 2) * 2);
    } else {
#pragma unroll
        for (int p = 0; p < 4; p++) w.v[p] = lds128(grp + p * 512 + lane * 16);
        w.s = lds32(grp + 2048 + lane * 4);
    }
}

template <int TYPE>
__device__ __forceinline__ float dot_w(const WRegs& w, int lane, const XRegs& x) {
    if constexpr (TYPE == T_Q4_K) {
        const int j = lane & 3; const uint4& hdr = w.v[0];
        int s0 = 0, s1 = 0, t0 = 0, t1 = 0;                        // four independent dp4a chains (depth 4 instead of 8)
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t a = u4_get(w.v[1], i), v = u4_get(w.v[2], i);
            s0 = dp4a_ss((int)(a & 0x0F0F0F0Fu), (int)u4_get(x.p[0], i), s0);
            s1 = dp4a_us(a & 0xF0F0F0F0u, (int)u4_get(x.p[2], i), s1);
            t0 = dp4a_ss((int)(v & 0x0F0F0F0Fu), (int)u4_get(x.p[1], i), t0);
            t1 = dp4a_us(v & 0xF0F0F0F0u, (int)u4_get(x.p[3], i), t1);
        }
        s0 += t0; s1 += t1;
        s1 >>= 4;                                                  // exact: every term is a multiple of 16
        int sc0, sc1, m0, m1; k4_scales(hdr, j, sc0, sc1, m0, m1);
        const int isum = sc0 * s0 + sc1 * s1;
        const int msum = m0 * (x.bs[0] + x.bs[1]) + m1 * (x.bs[2] + x.bs[3]);
        const float2 dd = __half22float2(*(const __half2*)&hdr.x);
        return x.dx * (dd.x * (float)isum - dd.y * (float)msum);
    } else if constexpr (TYPE == T_Q5_K) {
        const int j = lane & 3; const uint4& hdr = w.v[0];
        int s0 = 0, s1 = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            uint32_t q = u4_get(w.v[3], i), h = u4_get(w.v[1], i) >> (2 * j);
            s0 = dp4a_ss((int)((q & 0x0F0F0F0Fu) | ((h & 0x01010101u) << 4)), (int)u4_get(x.p[0], i), s0);
            s1 = dp4a_ss((int)(((q >> 4) & 0x0F0F0F0Fu) | ((h & 0x02020202u) << 3)), (int)u4_get(x.p[2], i), s1);
            q = u4_get(w.v[4], i); h = u4_get(w.v[2], i) >> (2 * j);
            s0 = dp4a_ss((int)((q & 0x0F0F0F0Fu) | ((h & 0x01010101u) << 4)), (int)u4_get(x.p[1], i), s0);
            s1 = dp4a_ss((int)(((q >> 4) & 0x0F0F0F0Fu) | ((h & 0x02020202u) << 3)), (int)u4_get(x.p[3], i), s1);
        }
        int sc0, sc1, m0, m1; k4_scales(hdr, j, sc0, sc1, m0, m1);
        const int isum = sc0 * s0 + sc1 * s1;
        const int msum = m0 * (x.bs[0] + x.bs[1]) + m1 * (x.bs[2] + x.bs[3]);
        const float2 dd = __half22float2(*(const __half2*)&hdr.x);
        return x.dx * (dd.x * (float)isum - dd.y * (float)msum);
    } else if constexpr (TYPE == T_Q6_K) {
        const uint32_t sc4 = w.v[3].x;
        const float d = __half2float(__ushort_as_half((unsigned short)w.s));
        int s[4] = {0, 0, 0, 0};
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t a = u4_get(w.v[0], i), b = u4_get(w.v[1], i), h = u4_get(w.v[2], i);
            s[0] = dp4a_ss((int)((a & 0x0F0F0F0Fu) | ((h << 4) & 0x30303030u)), (int)u4_get(x.p[0], i), s[0]);
            s[1] = dp4a_ss((int)((b & 0x0F0F0F0Fu) | ((h << 2) & 0x30303030u)), (int)u4_get(x.p[1], i), s[1]);
            s[2] = dp4a_ss((int)(((a >> 4) & 0x0F0F0F0Fu) | (h & 0x30303030u)), (int)u4_get(x.p[2], i), s[2]);
            s[3] = dp4a_ss((int)(((b >> 4) & 0x0F0F0F0Fu) | ((h >> 2) & 0x30303030u)), (int)u4_get(x.p[3], i), s[3]);
        }
        int tot = 0;
#pragma unroll
        for (int g = 0; g < 4; g++) tot += (int)(int8_t)(sc4 >> (8 * g)) * (s[g] - 32 * x.bs[g]);
        return x.dx * d * (float)tot;
    } else {
        int s0 = 0, s1 = 0;
#pragma unroll
        for (int p = 0; p < 4; p++) {
#pragma unroll
            for (int i = 0; i < 4; i++) {
                if (p < 2) s0 = dp4a_ss((int)u4_get(w.v[p], i), (int)u4_get(x.p[p], i), s0);
                else s1 = dp4a_ss((int)u4_get(w.v[p], i), (int)u4_get(x.p[p], i), s1);
            }
        }
        const float2 dd = __half22float2(*(const __half2*)&w.s);
        return x.dx * (dd.x * (float)s0 + dd.y * (float)s1);
    }
}

template <int TYPE>
__device__ __forceinline__ float lane_dot(const uint8_t* grp, int lane, const XRegs& x) { WRegs w; load_w<TYPE>(grp, lane, w); return dot_w<TYPE>(w, lane, x); }

}  // namespace b200
