// sampler.cuh — K8: temperature / top-k / top-p / multinomial sampling over the logits of one
// sequence (greedy argmax is fused into the lm_head epilogue in decode.cuh).  Order of operations
// follows llama.cpp's default chain [external]: top-k cut -> top-p nucleus on T=1 probabilities ->
// temperature -> categorical draw.  Deterministic for a given (seed, step): candidates are totally
// ordered by (logit desc, index asc) and the draw is Philox4x32-10(seed, step).
// Limit: at most SAMPLER_CAND candidates survive the top-k stage (top_k = 0 means SAMPLER_CAND).
#pragma once
#include <stdint.h>

namespace b200 {

constexpr int SAMPLER_THREADS = 1024;
constexpr int SAMPLER_CAND = 1024;
constexpr int SAMPLER_BUF = 2048;

__host__ __device__ inline void philox4x32_10(uint64_t seed, uint64_t ctr, uint32_t out[4]) {
    uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32), c2 = 0, c3 = 0, k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3; k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

__device__ __forceinline__ uint32_t float_key(float f) { uint32_t b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }

// Per-row sampling request of the batched path (device copy of the b200_sampling fields the draw needs + the draw counter).
struct RowSampling { float temperature, top_p; int top_k, pad; uint64_t seed, step; };

// Logits of one sequence.  A tensor-parallel engine holds them in `parts` slices of V_l vocabulary entries each
// ([part][row][ldv], see Model::dense_rows); a single GPU has one part.
struct LogitView {
    const float* base; int V, V_l; size_t part_stride;
    __device__ __forceinline__ float at(int i) const { return V_l >= V ? base[i] : base[(size_t)(i / V_l) * part_stride + (i % V_l)]; }
};

// inclusive prefix sum of one value per thread over the block (fixed order: warp shuffles, then the warp totals)
__device__ __forceinline__ float block_scan(float v, float* warp_tot /* [32] shared */, float& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const float t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
    __syncthreads();                                   // warp_tot free (previous scan's readers done)
    if (lane == 31) warp_tot[warp] = v;
    __syncthreads();
    float pre = 0.f, tot = 0.f;
#pragma unroll
    for (int w = 0; w < SAMPLER_THREADS / 32; w++) { const float t = warp_tot[w]; if (w < warp) pre += t; tot += t; }
    total = tot;
    return v + pre;
}

// One block samples one token.  Shared by the batch-1 kernel and the per-row kernel of the batched path, so a
// request's tokens do not depend on which path (or which batch) served it.
__device__ __forceinline__ int sample_block(const LogitView& lg, int top_k, float top_p, float temperature, uint64_t seed, uint64_t step) {
    __shared__ unsigned hist[256];
    __shared__ float cv[SAMPLER_BUF]; __shared__ int ci[SAMPLER_BUF];
    __shared__ float warp_tot[32];
    __shared__ unsigned n_cand; __shared__ uint32_t sh_prefix; __shared__ int sh_remaining, sh_first;
    const int tid = threadIdx.x, V = lg.V;
    const int K = (top_k > 0 && top_k < SAMPLER_CAND) ? min(top_k, V) : min(SAMPLER_CAND, V);
    // ---- radix select: key of the K-th largest logit
    uint32_t prefix = 0, mask = 0; int remaining = K;
    for (int shift = 24; shift >= 0; shift -= 8) {
        if (tid < 256) hist[tid] = 0;
        __syncthreads();
        for (int i = tid; i < V; i += SAMPLER_THREADS) { uint32_t k = float_key(lg.at(i)); if ((k & mask) == prefix) atomicAdd(&hist[(k >> shift) & 255], 1u); }
        __syncthreads();
        if (tid == 0) {
            int rem = remaining, b = 255;
            for (; b > 0; b--) { if ((int)hist[b] >= rem) break; rem -= hist[b]; }
            sh_prefix = prefix | ((uint32_t)b << shift); sh_remaining = rem;
        }
        __syncthreads();
        prefix = sh_prefix; remaining = sh_remaining; mask |= 0xFFu << shift;
        __syncthreads();
    }
    // ---- gather everything >= the K-th key, pad, sort by (logit desc, index asc)
    if (tid == 0) n_cand = 0;
    __syncthreads();
    for (int i = tid; i < V; i += SAMPLER_THREADS) {
        float v = lg.at(i);
        if (float_key(v) >= prefix) { unsigned s = atomicAdd(&n_cand, 1u); if (s < SAMPLER_BUF) { cv[s] = v; ci[s] = i; } }
    }
    __syncthreads();
    const int n = min((int)n_cand, SAMPLER_BUF);
    for (int i = tid; i < SAMPLER_BUF; i += SAMPLER_THREADS) if (i >= n) { cv[i] = -INFINITY; ci[i] = 0x7fffffff; }
    __syncthreads();
    for (int k = 2; k <= SAMPLER_BUF; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < SAMPLER_BUF; i += SAMPLER_THREADS) {
                int p = i ^ j;
                if (p > i) {
                    bool before_ip = cv[i] > cv[p] || (cv[i] == cv[p] && ci[i] < ci[p]);     // i should precede p
                    bool asc_block = (i & k) != 0;                                           // reversed sub-sequence
                    if (before_ip == asc_block) { float tv = cv[i]; cv[i] = cv[p]; cv[p] = tv; int ti = ci[i]; ci[i] = ci[p]; ci[p] = ti; }
                }
            }
            __syncthreads();
        }
    // ---- nucleus at T=1, then the draw at temperature: candidate i belongs to thread i (K <= 1024 = block size);
    //      cumulative sums by block-wide scans, cut points by an atomic minimum over the threads that satisfy them
    const int kk = min(K, n);
    const float mx = cv[0];
    int keep = kk;
    if (top_p < 1.f) {
        float tot; const float cum = block_scan(tid < kk ? expf(cv[tid] - mx) : 0.f, warp_tot, tot);
        if (tid == 0) sh_first = kk - 1;
        __syncthreads();
        if (tid < kk && cum >= top_p * tot) atomicMin(&sh_first, tid);
        __syncthreads();
        keep = sh_first + 1;
        __syncthreads();
    }
    const float invT = 1.f / temperature;
    float tot2; const float cum2 = block_scan(tid < keep ? expf((cv[tid] - mx) * invT) : 0.f, warp_tot, tot2);
    uint32_t r[4]; philox4x32_10(seed, step, r);
    const float u = (float)(r[0] >> 8) * (1.0f / 16777216.0f) * tot2;
    if (tid == 0) sh_first = keep - 1;
    __syncthreads();
    if (tid < keep && cum2 > u) atomicMin(&sh_first, tid);
    __syncthreads();
    return ci[sh_first];
}

__global__ void __launch_bounds__(SAMPLER_THREADS, 1)
sample_kernel(const float* __restrict__ logits, int V, int top_k, float top_p, float temperature, uint64_t seed, uint64_t step,
              int* __restrict__ token_out, int* __restrict__ out_slot) {
    const LogitView lg{logits, V, V, 0};
    const int t = sample_block(lg, top_k, top_p, temperature, seed, step);
    if (threadIdx.x == 0) { *token_out = t; if (out_slot) *out_slot = t; }
}

// Batched decode: block b draws row b's token when that row asked for sampling (temperature > 0); greedy rows keep the
// argmax already in next[b].  Parameters come from device memory so the launch is identical from step to step (CUDA graph).
__global__ void __launch_bounds__(SAMPLER_THREADS, 1)
sample_rows_kernel(const float* __restrict__ logits, int V, int V_l, int ldv, size_t part_stride, const RowSampling* __restrict__ rs, int* __restrict__ next) {
    const RowSampling s = rs[blockIdx.x];
    if (!(s.temperature > 0.f)) return;
    const LogitView lg{logits + (size_t)blockIdx.x * ldv, V, V_l, part_stride};
    const int t = sample_block(lg, s.top_k, s.top_p <= 0.f ? 1.f : s.top_p, s.temperature, s.seed, s.step);
    if (threadIdx.x == 0) next[blockIdx.x] = t;
}

}  // namespace b200
