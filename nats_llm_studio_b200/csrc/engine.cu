// engine.cu — host side of the engine: HBM-resident model registry, GGUF upload + on-GPU repack,
// decode-schedule construction, kernel launches, and the kernel-level C ABI entry points.
// Replaces what /root/reference/nats_llm_studio.go delegates to LM Studio (Chat :158-179,
// PullModel :46-59, ListModels :136-156, DeleteModel :99-133).
#include "engine.hpp"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "decode.cuh"
#include "prefill.cuh"
#include "qgemm.cuh"
#include "attn_tc.cuh"
#include "sampler.cuh"

namespace b200 {

thread_local std::string g_last_error;
int fail(int code, const std::string& msg) {
    g_last_error = msg;
    static const bool log = getenv("B200_LOG_ERRORS") != nullptr;       // errors raised on the scheduler thread are otherwise only visible as event codes
    if (log) fprintf(stderr, "[b200] error %d: %s\n", code, msg.c_str());
    return code;
}

#define CK(expr)                                                                                              \
    do {                                                                                                      \
        cudaError_t e_ = (expr);                                                                              \
        if (e_ != cudaSuccess) return fail(B200_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e_)); \
    } while (0)

Engine& engine() { static Engine e; return e; }

// ------------------------------------------------------------------------------------ tensor parallelism: NCCL, bound at run time
// (no link-time dependency: the library loads on a box without NCCL; one process per GPU)
#include <dlfcn.h>
namespace {
struct NcclId { char internal[128]; };
struct NcclApi {
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
    int (*CommDestroy)(void*) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok = false;
};
NcclApi& nccl() {
    static NcclApi a; static bool tried = false;
    if (!tried) {
        tried = true;
        void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (h) {
            a.GetUniqueId = (decltype(a.GetUniqueId))dlsym(h, "ncclGetUniqueId"); a.CommInitRank = (decltype(a.CommInitRank))dlsym(h, "ncclCommInitRank");
            a.AllReduce = (decltype(a.AllReduce))dlsym(h, "ncclAllReduce"); a.AllGather = (decltype(a.AllGather))dlsym(h, "ncclAllGather");
            a.CommDestroy = (decltype(a.CommDestroy))dlsym(h, "ncclCommDestroy"); a.GetErrorString = (decltype(a.GetErrorString))dlsym(h, "ncclGetErrorString");
            a.ok = a.GetUniqueId && a.CommInitRank && a.AllReduce && a.AllGather && a.CommDestroy && a.GetErrorString;
        }
    }
    return a;
}
int nccl_fail(const char* what, int rc) { return fail(B200_ERR_CUDA, std::string(what) + ": " + (nccl().GetErrorString ? nccl().GetErrorString(rc) : "NCCL error")); }
}  // namespace

int Engine::tp_unique_id(void* out, size_t cap) {
    if (cap < sizeof(NcclId)) return fail(B200_ERR_INVALID, "the NCCL id needs 128 bytes");
    if (!nccl().ok) return fail(B200_ERR_UNSUPPORTED, "libnccl.so.2 not found: tensor parallelism is unavailable");
    if (int rc = nccl().GetUniqueId((NcclId*)out)) return nccl_fail("ncclGetUniqueId", rc);
    return B200_OK;
}
int Engine::tp_init(int rank, int world, const void* id, size_t n) {
    if (int rc = ensure()) return rc;
    if (world < 1 || rank < 0 || rank >= world || (world > 1 && (!id || n < sizeof(NcclId)))) return fail(B200_ERR_INVALID, "bad tensor-parallel rank/world/id");
    if (nccl_comm) return fail(B200_ERR_STATE, "tensor parallelism is already initialised");
    if (world > 1) {
        if (!nccl().ok) return fail(B200_ERR_UNSUPPORTED, "libnccl.so.2 not found: tensor parallelism is unavailable");
        NcclId nid; memcpy(&nid, id, sizeof nid);
        if (int rc = nccl().CommInitRank(&nccl_comm, world, nid, rank)) return nccl_fail("ncclCommInitRank", rc);
    }
    tp_rank = rank; tp_world = world;
    return B200_OK;
}
int Engine::all_reduce_sum(float* buf, size_t n) {
    if (tp_world == 1) return B200_OK;
    if (int rc = nccl().AllReduce(buf, buf, n, 7 /*ncclFloat32*/, 0 /*ncclSum*/, nccl_comm, stream)) return nccl_fail("ncclAllReduce", rc);
    return B200_OK;
}
int Engine::all_gather(const float* src, float* dst, size_t n_per_rank) {
    if (tp_world == 1) { if (src != dst) CK(cudaMemcpyAsync(dst, src, n_per_rank * 4, cudaMemcpyDeviceToDevice, stream)); return B200_OK; }
    if (int rc = nccl().AllGather(src, dst, n_per_rank, 7, nccl_comm, stream)) return nccl_fail("ncclAllGather", rc);
    return B200_OK;
}

// Which part of a GGUF tensor rank `rank` of `world` keeps.  Column-parallel tensors (attn_q/k/v by
// heads, ffn_gate/up by FFN columns, output by vocabulary rows) keep a contiguous row range; row-parallel ones
// (attn_output, ffn_down) keep every row and a K window, which must start and end on 256-element block boundaries.
int tp_shard_window(const char* kind, int64_t N, int K, int rank, int world, int64_t* row0, int64_t* nrows, int* col0, int* ncols) {
    if (!kind || world < 1 || rank < 0 || rank >= world) return fail(B200_ERR_INVALID, "bad tensor-parallel rank/world");
    const std::string k = kind;
    const bool by_rows = k == "attn_q" || k == "attn_k" || k == "attn_v" || k == "ffn_gate" || k == "ffn_up" || k == "output";
    const bool by_cols = k == "attn_output" || k == "ffn_down";
    if (!by_rows && !by_cols) return fail(B200_ERR_INVALID, "tensor kind '" + k + "' is not sharded (norms and token_embd are replicated)");
    if (by_rows) {
        if (N % world) return fail(B200_ERR_UNSUPPORTED, k + ": " + std::to_string(N) + " rows do not divide over " + std::to_string(world) + " ranks");
        *nrows = N / world; *row0 = rank * *nrows; *col0 = 0; *ncols = K;
    } else {
        if (K % (256 * world)) return fail(B200_ERR_UNSUPPORTED, k + ": K " + std::to_string(K) + " does not divide over " + std::to_string(world) + " ranks in 256-element blocks");
        *ncols = K / world; *col0 = rank * *ncols; *row0 = 0; *nrows = N;
    }
    return B200_OK;
}

// ------------------------------------------------------------------------------------ init
// ------------------------------------------------------------------------------------ decode kernel variants
// The persistent kernel is instantiated for the model shapes that matter and once with everything; a model gets the
// smallest variant covering its head size, weight types and tensor parallelism (DK_* bits, stream.cuh).
struct DecodeVariant { int cfg; const void* fn; };
static const DecodeVariant g_decode_variants[] = {
    {DK_HD128 | DK_Q4K | DK_Q6K, (const void*)decode_kernel_t<DK_HD128 | DK_Q4K | DK_Q6K>},                 // Llama-3 Q4_K_M files
    {DK_HD128 | DK_Q80, (const void*)decode_kernel_t<DK_HD128 | DK_Q80>},                                     // Q8_0 files
    {DK_HD128 | DK_TP | DK_Q4K | DK_Q6K, (const void*)decode_kernel_t<DK_HD128 | DK_TP | DK_Q4K | DK_Q6K>},  // tensor-parallel shards of Q4_K_M files
    {DK_HD128 | DK_HD64 | DK_TYPES, (const void*)decode_kernel_t<DK_HD128 | DK_HD64 | DK_TYPES>},            // any single-GPU model
    {DK_HD128 | DK_HD64 | DK_TP | DK_TYPES, (const void*)decode_kernel_t<DK_HD128 | DK_HD64 | DK_TP | DK_TYPES>},   // any model
    {DK_HD128 | DK_Q4K | DK_Q6K | DK_TRACE, (const void*)decode_kernel_t<DK_HD128 | DK_Q4K | DK_Q6K | DK_TRACE>},  // B200_TRACE runs of the bench model
    {DK_ALL, (const void*)decode_kernel_t<DK_ALL>},
};
static int decode_variant_count() { return (int)(sizeof(g_decode_variants) / sizeof(g_decode_variants[0])); }
static const void* decode_variant_fn(int i) { return g_decode_variants[i].fn; }
static const void* decode_variant_for(int need) {
    if (getenv("B200_DECODE_GENERIC")) return g_decode_variants[decode_variant_count() - 1].fn;      // A/B measurements
    if (const char* e = getenv("B200_DECODE_VARIANT")) { const int i = atoi(e); if (i >= 0 && i < decode_variant_count() && (g_decode_variants[i].cfg & need) == need) return g_decode_variants[i].fn; }
    for (const DecodeVariant& v : g_decode_variants) if ((v.cfg & need) == need) return v.fn;         // ordered: specific first
    return g_decode_variants[decode_variant_count() - 1].fn;
}
static int dk_type_bit(int type) { return type == T_Q4_K ? DK_Q4K : type == T_Q5_K ? DK_Q5K : type == T_Q6_K ? DK_Q6K : DK_Q80; }

int Engine::init(const b200_config* cfg) {
    std::lock_guard<std::mutex> lk(mu);
    if (ready) return B200_OK;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(B200_ERR_CUDA, std::string("no CUDA device: the engine has no CPU fallback (") + cudaGetErrorString(e) + ")");
    device = cfg ? cfg->device : 0;
    max_ctx = cfg && cfg->max_ctx > 0 ? cfg->max_ctx : 4096;
    max_seqs = cfg && cfg->max_seqs > 0 ? cfg->max_seqs : 1;
    kv_pages = cfg ? (int)cfg->kv_pages : 0;
    // the batch-1 kernel caches a sequence's block table in shared memory (BTAB_MAX pages of 16 positions)
    if (max_ctx > 16 * BTAB_MAX) return fail(B200_ERR_UNSUPPORTED, "max_ctx " + std::to_string(max_ctx) + " exceeds the supported " + std::to_string(16 * BTAB_MAX) + " positions per sequence");
    if (device < 0 || device >= 64) return fail(B200_ERR_INVALID, "device ordinal out of range");
    CK(cudaSetDevice(device));
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, device));
    if (p.major < 10) return fail(B200_ERR_UNSUPPORTED, "this library is built for sm_100a (B200) only");
    n_sm = p.multiProcessorCount; smem_per_sm = (int)p.sharedMemPerMultiprocessor; smem_optin = (int)p.sharedMemPerBlockOptin;
    CK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&aux_stream, cudaStreamNonBlocking));
    for (int i = 0; i < decode_variant_count(); i++) {
#ifdef B200_WTRACE
        CK(cudaFuncSetAttribute(decode_variant_fn(i), cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin - 1024));   // the variant has a static shared array
#else
        CK(cudaFuncSetAttribute(decode_variant_fn(i), cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin));
#endif
    }
    ready = true;
    return B200_OK;
}

void Engine::shutdown() {
    std::lock_guard<std::mutex> lk(mu);
    if (!ready) return;
    cudaSetDevice(device); cudaDeviceSynchronize();
    cudaStreamDestroy(stream); cudaStreamDestroy(copy_stream); cudaStreamDestroy(aux_stream); stream = copy_stream = aux_stream = nullptr;
    cudaFree(qg_ws); qg_ws = nullptr;
    if (nccl_comm) { nccl().CommDestroy(nccl_comm); nccl_comm = nullptr; }
    tp_rank = 0; tp_world = 1;
    ready = false;
}

int Engine::ensure() {
    if (ready) { cudaSetDevice(device); return B200_OK; }
    return init(nullptr);
}

// ------------------------------------------------------------------------------------ launch geometry
struct LaunchCfg { int grid; uint32_t ring_bytes; int max_grp; size_t smem; const void* kernel = nullptr; };

static int launch_cfg(Engine& E, int max_grp, int hown_cap, int max_segs, int need, LaunchCfg& lc) {
    // one persistent CTA per SM owning (almost) all of its shared memory
    max_grp = std::max(max_grp, (int)(((CONSUMER_WARPS + 1) * 130 + 64) * 4 / 2048 + 1));   // the int8 activation buffer doubles as attention scratch
    if (max_grp > XG_MAX) return fail(B200_ERR_UNSUPPORTED, "activation vector longer than " + std::to_string(XG_MAX * 2048) + " elements");
    size_t per_cta = (size_t)E.smem_optin;
    size_t fixed = SmemLayout::RING + smem_tail_bytes(max_grp, hown_cap, max_segs);
    // the ring allocator never splits a stage across the wrap point, so a stage plus the gap it may leave behind
    // must fit: ring >= 2 x the largest stage (otherwise the producer can wait on space that cannot exist)
    if (fixed + 2 * MAX_STAGE_BYTES + 256 > per_cta) return fail(B200_ERR_UNSUPPORTED, "activation vector does not fit in shared memory");
    lc.grid = E.n_sm;
    lc.ring_bytes = (uint32_t)((per_cta - fixed - 128) / 128 * 128);
    // Not all of it: shared memory and L1 share one 228 KB array and the split moves in steps (… 132, 164, 196, 228 KB of
    // shared memory, counting 1 KB the system reserves per CTA).  The kernel's register spills and exchange polls live in
    // L1, so the ring takes what is left of the 196 KB step (32 KB of L1): measured on the 8B model 587 tok/s, against 566
    // at the 164 KB step (64 KB of L1, ring 132 KB) and 555 one byte over the 196 KB step (no L1 at all).  B200_RING_KB overrides.
    uint32_t want = fixed + 2u * MAX_STAGE_BYTES < 195u * 1024u ? (uint32_t)(195u * 1024u - fixed) / 128 * 128 : 0u;      // 0: keep all of it
    if (const char* rk = getenv("B200_RING_KB")) want = (uint32_t)atoi(rk) * 1024u;
    if (want >= 2u * MAX_STAGE_BYTES && want < lc.ring_bytes) lc.ring_bytes = want / 128 * 128;
    lc.max_grp = max_grp; lc.smem = lc.ring_bytes + fixed;
    lc.kernel = decode_variant_for(need);
    int occ = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, lc.kernel, CTA_THREADS, lc.smem));
    if (occ < 1) return fail(B200_ERR_CUDA, "decode kernel does not fit on an SM");
    return B200_OK;
}

static int launch_decode(Engine& E, const DecodeParams& P, const DecodeRun& R, const LaunchCfg& lc) {
    void* args[] = {(void*)&P, (void*)&R, (void*)&lc.ring_bytes, (void*)&lc.max_grp};
    CK(cudaLaunchCooperativeKernel(lc.kernel, dim3(lc.grid), dim3(CTA_THREADS), args, lc.smem, E.stream));
    E.launches++;
    return B200_OK;
}

// ------------------------------------------------------------------------------------ device helpers
template <typename T>
static int dalloc(T** p, size_t n) {
    cudaError_t e = cudaMalloc((void**)p, n * sizeof(T));
    if (e != cudaSuccess) return fail(B200_ERR_NOMEM, std::string("cudaMalloc of ") + std::to_string(n * sizeof(T)) + " bytes: " + cudaGetErrorString(e));
    return B200_OK;
}

// repack `raw` (device, ggml row-major blocks) into the engine layout at `out`
// (N rows written, of which the first n_src come from raw and the rest are zero padding)
static int repack(Engine& E, const uint8_t* raw, uint8_t* out, int type, int64_t N, int64_t n_src, int K, int interleave_parity = -1, size_t src_stride = 0, cudaStream_t st = nullptr) {
    if (!st) st = E.stream;
    long long chunks = (long long)N * n_groups(K) * group_bytes(type) / 16;
    int blocks = (int)std::min<long long>((chunks + 255) / 256, 148LL * 32);
    if (blocks < 1) blocks = 1;
    repack_kernel<<<blocks, 256, 0, st>>>(raw, out, type, N, n_src, K, interleave_parity, src_stride);
    CK(cudaGetLastError());
    E.launches++;
    return B200_OK;
}

// ------------------------------------------------------------------------------------ Model
Model::~Model() { release(); }

// Frees every device allocation and forgets every pointer and "already built" flag derived from them: a request that
// still holds the model (a queued chat, the scheduler's active sequences) finds committed == false under run_mu and
// fails with B200_ERR_STATE instead of launching kernels on freed memory.  Caller holds run_mu (or is the destructor).
void Model::release_staging() {
    for (auto& sl : stage_live) cudaEventDestroy(sl.freed);
    stage_live.clear();
    if (d_stage) { cudaFree(d_stage); d_stage = nullptr; }
    t_got.clear(); t_ring.clear(); t_order.clear();
}

void Model::release() {
    if (engine().ready) { cudaSetDevice(engine().device); cudaStreamSynchronize(engine().stream); }
    for (auto& g : step_graph) if (g) { cudaGraphExecDestroy(g); g = nullptr; }
    step_graph.clear(); step_seen.clear(); step_launches.clear(); graph_failed = false;
    for (int r = 0; r < 8; r++) { if (tp_peer[r] && tp_peer[r] != tp_buf) cudaIpcCloseMemHandle(tp_peer[r]); tp_peer[r] = nullptr; }
    tp_buf = nullptr; tp_error = nullptr;
    for (void* p : owned) cudaFree(p);
    owned.clear();
    release_staging();
    committed = false;
    tok_embd = Matrix{}; out_w = Matrix{}; out_norm = nullptr; inv_freq = nullptr; layers.clear();
    d_phases = nullptr; n_phases = 0; d_logits = nullptr; d_ll16 = nullptr; d_ll32 = nullptr;
    d_state = d_out_tokens = d_forced = d_block_table = nullptr; d_kv = nullptr; n_pages = 0; cur_pos = 0;
    pf_H = pf_Y = nullptr; pf_X = pf_Q = pf_ACT = pf_W = nullptr; pf_toks = nullptr; pf_T = nullptr;
    bt_logits = nullptr; bt_next = nullptr; bt_samp = nullptr; b_out = nullptr; tmap_q = tmap_kv = tmap_qd = nullptr;
    bf16_ready = bf16_tried = tile_ready = false;
    hbm_bytes = 0; weight_bytes = 0; uploaded = 0; t_done.clear(); rope_ff.clear();
}

template <typename T>
int Model::own(T** p, size_t n) { int rc = dalloc(p, n); if (rc == B200_OK) { owned.push_back(*p); hbm_bytes += n * sizeof(T); } return rc; }

int Model::open(const std::string& id_, const uint8_t* head, size_t n) {
    id = id_;
    std::string err = gguf_parse(head, n, meta);
    if (err == "need more") return fail(B200_ERR_MORE, "GGUF header longer than the supplied prefix");
    if (!err.empty()) return fail(B200_ERR_INVALID, "GGUF: " + err);
    arch = meta.str("general.architecture");
    if (arch != "llama" && arch != "granite") return fail(B200_ERR_UNSUPPORTED, "unsupported architecture '" + arch + "' (llama, granite)");
    auto an = [&](const char* k, double d) { return meta.num(arch + "." + k, d); };
    n_layer = (int)an("block_count", 0); n_embd = (int)an("embedding_length", 0); n_head = (int)an("attention.head_count", 0);
    n_head_kv = (int)an("attention.head_count_kv", n_head); n_ff = (int)an("feed_forward_length", 0);
    n_ctx_train = (int)an("context_length", 4096);
    rms_eps = (float)an("attention.layer_norm_rms_epsilon", 1e-5); rope_base = (float)an("rope.freq_base", 10000.0);
    head_dim = n_head ? (int)an("attention.key_length", n_embd / n_head) : 0;
    n_rot = (int)an("rope.dimension_count", head_dim);
    emb_scale = (float)an("embedding_scale", 0); res_scale = (float)an("residual_scale", 0);
    attn_scale = (float)an("attention.scale", 0); logit_scale = (float)an("logit_scale", 0);
    const GgufTensor* te = meta.tensor("token_embd.weight");
    if (!te || !n_layer || !n_embd || !n_head || !n_ff) return fail(B200_ERR_INVALID, "GGUF: missing hyper-parameters or token_embd.weight");
    n_vocab = (int)te->ne[1];
    if (head_dim != 64 && head_dim != 128) return fail(B200_ERR_UNSUPPORTED, "head_dim must be 64 or 128");
    if (n_embd % 256 || n_ff % 256 || (n_head * head_dim) % 256) return fail(B200_ERR_UNSUPPORTED, "dimensions must be multiples of 256");
    if (n_head % n_head_kv) return fail(B200_ERR_INVALID, "head_count not a multiple of head_count_kv");
    bos = (int)meta.num("tokenizer.ggml.bos_token_id", -1); eos = (int)meta.num("tokenizer.ggml.eos_token_id", -1);
    if (auto v = meta.find("tokenizer.ggml.tokens")) vocab = v->arr_str;
    if (auto v = meta.find("tokenizer.ggml.token_type")) { tok_type.assign(v->arr_num.size(), 0); for (size_t i = 0; i < v->arr_num.size(); i++) tok_type[i] = (int)v->arr_num[i]; }
    build_tokenizer();
    Engine& E = engine();
    if (int rc = E.ensure()) return rc;
    // tensor parallelism: this rank keeps heads / FFN columns / vocabulary rows divided by the world size (tp_shard_window)
    const int W = E.tp_world;
    tp = W > 1; g_head = n_head; g_head_kv = n_head_kv; g_ff = n_ff;
    if (tp) {
        if (n_head_kv % W || n_head % W || n_ff % (256 * W) || (n_head * head_dim / W) % 256 || n_vocab % (2 * W))
            return fail(B200_ERR_UNSUPPORTED, "model dimensions do not divide over " + std::to_string(W) + " tensor-parallel ranks");
        n_head /= W; n_head_kv /= W; n_ff /= W; n_vocab_l = n_vocab / W;
    } else n_vocab_l = n_vocab;
    layers.assign(n_layer, Layer{});
    // ---- staging: uploads land in a ring in HBM, tensor by tensor in file order; a tensor is repacked into its final layout
    // as soon as its last byte has arrived and its ring space is reused once that repack has run (K9 "behind the H2D
    // stream").  Peak HBM during a load = the resident weights + the ring (at least two of the largest tensor).
    t_order.resize(meta.tensors.size());
    for (size_t i = 0; i < t_order.size(); i++) t_order[i] = (int)i;
    std::sort(t_order.begin(), t_order.end(), [&](int x, int y) { return meta.tensors[x].offset < meta.tensors[y].offset; });
    uint64_t biggest = 0; for (auto& t : meta.tensors) biggest = std::max(biggest, t.nbytes);
    stage_bytes = std::min<uint64_t>(std::max<uint64_t>(2 * ((biggest + 255) & ~255ull) + 512, 1ull << 30), ((meta.data_bytes + 255) & ~255ull) + 512);
    if (const char* sb = getenv("B200_STAGE_MB")) stage_bytes = std::max<uint64_t>(2 * ((biggest + 255) & ~255ull) + 512, (uint64_t)atoll(sb) << 20);
    t_got.assign(meta.tensors.size(), 0); t_ring.assign(meta.tensors.size(), ~0ull); t_cursor = 0; stage_head = 0; next_off = 0;
    cudaError_t e = cudaMalloc((void**)&d_stage, stage_bytes);
    if (e != cudaSuccess) return fail(B200_ERR_NOMEM, std::string("staging cudaMalloc: ") + cudaGetErrorString(e));
    return B200_OK;
}

// ring space for one tensor (contiguous: the repack reads rows with a stride); blocks until earlier tenants were repacked
int Model::stage_alloc(uint64_t nbytes, uint64_t* off) {
    const uint64_t need = (nbytes + 255) & ~255ull;
    if (need > stage_bytes) return fail(B200_ERR_NOMEM, "tensor larger than the staging ring");
    if (stage_head + need > stage_bytes) stage_head = 0;
    const uint64_t b0 = stage_head, b1 = stage_head + need;
    while (!stage_live.empty() && stage_live.front().begin < b1 && b0 < stage_live.front().end) {
        StageSlot& f = stage_live.front();
        if (!f.recorded) return fail(B200_ERR_STATE, "staging ring too small for the tensors in flight");
        CK(cudaEventSynchronize(f.freed)); cudaEventDestroy(f.freed); stage_live.pop_front();
    }
    *off = b0; stage_head = b1;
    return B200_OK;
}

// Bytes [file_off, file_off + n) of the .gguf file.  Uploads must arrive in file order (a chunked Object Store read, a
// sequential file read).  The copy is queued on the copy stream; `ticket` (optional) receives an event the caller waits on
// before it reuses `bytes`, which therefore should be pinned memory.
int Model::upload_async(uint64_t file_off, const void* bytes, size_t n, cudaEvent_t* ticket) {
    if (committed || !d_stage) return fail(B200_ERR_STATE, "upload after commit");
    Engine& E = engine(); E.ensure();
    uint64_t lo = std::max<uint64_t>(file_off, meta.data_offset), hi = std::min<uint64_t>(file_off + n, meta.data_offset + meta.data_bytes);
    if (lo < hi) {
        lo -= meta.data_offset; hi -= meta.data_offset;                          // data-section offsets
        if (lo > next_off) return fail(B200_ERR_STATE, "uploads must arrive in file order (got data offset " + std::to_string(lo) + ", expected " + std::to_string(next_off) + ")");
        if (hi > next_off) {
            lo = std::max(lo, next_off);
            const uint8_t* src = (const uint8_t*)bytes + (lo + meta.data_offset - file_off);
            while (t_cursor < t_order.size()) {
                const int ti = t_order[t_cursor]; const GgufTensor& t = meta.tensors[(size_t)ti];
                if (hi <= t.offset) break;                                       // only alignment padding before the next tensor
                const uint64_t a = std::max(lo, t.offset), b = std::min(hi, t.offset + t.nbytes);
                if (a < b) {
                    if (t_ring[(size_t)ti] == ~0ull) {
                        uint64_t off; if (int rc = stage_alloc(t.nbytes, &off)) return rc;
                        t_ring[(size_t)ti] = off;
                        StageSlot sl{}; sl.begin = off; sl.end = off + ((t.nbytes + 255) & ~255ull); sl.recorded = false; sl.tensor = ti;
                        CK(cudaEventCreateWithFlags(&sl.freed, cudaEventDisableTiming)); stage_live.push_back(sl);
                    }
                    CK(cudaMemcpyAsync(d_stage + t_ring[(size_t)ti] + (a - t.offset), src + (a - lo), b - a, cudaMemcpyHostToDevice, E.copy_stream));
                    t_got[(size_t)ti] += b - a; uploaded += b - a;
                }
                if (t_got[(size_t)ti] < t.nbytes) break;                        // the rest of this tensor comes with the next upload
                // complete: repack behind the copy, then release the ring space
                cudaEvent_t landed; CK(cudaEventCreateWithFlags(&landed, cudaEventDisableTiming));
                CK(cudaEventRecord(landed, E.copy_stream)); CK(cudaStreamWaitEvent(E.stream, landed, 0)); cudaEventDestroy(landed);
                if (t.nbytes) { if (int rc = process_tensor(t, d_stage + t_ring[(size_t)ti])) return rc; }
                for (auto& sl : stage_live) if (sl.tensor == ti) { CK(cudaEventRecord(sl.freed, E.stream)); sl.recorded = true; }
                t_cursor++;
            }
            next_off = std::max(next_off, hi);
        }
    }
    if (ticket) { CK(cudaEventCreateWithFlags(ticket, cudaEventDisableTiming)); CK(cudaEventRecord(*ticket, E.copy_stream)); }
    return B200_OK;
}
int Model::upload(uint64_t file_off, const void* bytes, size_t n) {
    if (int rc = upload_async(file_off, bytes, n, nullptr)) return rc;
    CK(cudaStreamSynchronize(engine().copy_stream));                        // caller may reuse its buffer on return
    return B200_OK;
}

// One complete tensor (device bytes at `src`, ggml row-major blocks) -> its place in the model: this rank's shard of a
// matrix repacked into the engine layout, a norm vector copied, gate / up interleaved row-wise into one matrix.
int Model::process_tensor(const GgufTensor& t, const uint8_t* src) {
    Engine& E = engine(); const int W = E.tp_world, tr = E.tp_rank;
    const int D = n_embd, F = n_ff;
    const int64_t gQD = (int64_t)g_head * head_dim, gKVD = (int64_t)g_head_kv * head_dim;
    auto matrix = [&](const char* kind, int64_t N_full, int K_full, Matrix& out) -> int {      // kind == nullptr: the whole matrix
        if (t.ne[0] != K_full || t.ne[1] != N_full) return fail(B200_ERR_INVALID, "GGUF: tensor " + t.name + " has unexpected shape");
        if (!is_streamable(t.type)) return fail(B200_ERR_UNSUPPORTED, "tensor " + t.name + ": ggml type " + std::to_string(t.type) + " is not supported for matrices (Q4_K, Q5_K, Q6_K, Q8_0)");
        int64_t r0 = 0, nr = N_full; int c0 = 0, nc = K_full;
        if (kind) { if (int rc = tp_shard_window(kind, N_full, K_full, tr, W, &r0, &nr, &c0, &nc)) return rc; }
        if (c0 % 256 || nc % 256) return fail(B200_ERR_UNSUPPORTED, "tensor " + t.name + ": shard is not aligned to 256-element blocks");
        const int64_t Np = nr + (nr & 1);                                    // odd matrices (vocab) get one zero row
        out.type = t.type; out.N = Np; out.n_rows = nr; out.K = nc;
        if (int rc = own(&out.w, repacked_bytes(t.type, Np, nc))) return rc;
        const size_t rrb = raw_row_bytes(t.type, K_full);
        weight_bytes += (uint64_t)nr * raw_row_bytes(t.type, nc);
        return repack(E, src + (size_t)r0 * rrb + (size_t)(c0 / block_elems(t.type)) * block_bytes(t.type), out.w, t.type, Np, nr, nc, -1, rrb);
    };
    auto vector = [&](int64_t n, float** out) -> int {
        if (t.type != T_F32 || t.n_elems() != n) return fail(B200_ERR_UNSUPPORTED, "tensor " + t.name + " must be F32[" + std::to_string(n) + "]");
        if (int rc = own(out, (size_t)n)) return rc;
        weight_bytes += t.nbytes;
        CK(cudaMemcpyAsync(*out, src, (size_t)n * 4, cudaMemcpyDeviceToDevice, E.stream));
        return B200_OK;
    };
    int rc = B200_OK;
    if (t.name == "token_embd.weight") {
        rc = matrix(nullptr, n_vocab, D, tok_embd);                          // replicated on every rank
        if (!rc && tp && !meta.tensor("output.weight")) rc = matrix("output", n_vocab, D, out_w);      // tied lm_head: this rank's vocabulary rows
    } else if (t.name == "output.weight") rc = matrix(tp ? "output" : nullptr, n_vocab, D, out_w);
    else if (t.name == "output_norm.weight") rc = vector(D, &out_norm);
    else if (t.name == "rope_freqs.weight") {
        if (t.type == T_F32 && t.n_elems() == n_rot / 2) { rope_ff.assign((size_t)n_rot / 2, 1.f); CK(cudaMemcpyAsync(rope_ff.data(), src, rope_ff.size() * 4, cudaMemcpyDeviceToHost, E.stream)); CK(cudaStreamSynchronize(E.stream)); }
    } else if (t.name.compare(0, 4, "blk.") == 0) {
        const size_t dot = t.name.find('.', 4); if (dot == std::string::npos) return B200_OK;
        const int l = atoi(t.name.substr(4, dot - 4).c_str()); const std::string what = t.name.substr(dot + 1);
        if (l < 0 || l >= n_layer) return B200_OK;
        Layer& L = layers[(size_t)l];
        if (what == "attn_norm.weight") rc = vector(D, &L.attn_norm);
        else if (what == "ffn_norm.weight") rc = vector(D, &L.ffn_norm);
        else if (what == "attn_q.weight") rc = matrix("attn_q", gQD, D, L.wq);                 // column-parallel: by heads
        else if (what == "attn_k.weight") rc = matrix("attn_k", gKVD, D, L.wk);
        else if (what == "attn_v.weight") rc = matrix("attn_v", gKVD, D, L.wv);
        else if (what == "attn_output.weight") rc = matrix("attn_output", D, (int)gQD, L.wo);   // row-parallel: K range
        else if (what == "ffn_down.weight") rc = matrix("ffn_down", D, g_ff, L.wd);
        else if (what == "ffn_gate.weight" || what == "ffn_up.weight") {                        // rows interleaved (gate_i, up_i) in one matrix
            if (t.ne[0] != D || t.ne[1] != g_ff) return fail(B200_ERR_INVALID, "GGUF: " + t.name + " has unexpected shape");
            if (!is_streamable(t.type)) return fail(B200_ERR_UNSUPPORTED, t.name + ": unsupported ggml type");
            if (!L.wgu.w) {
                L.wgu.type = t.type; L.wgu.N = 2 * (int64_t)F; L.wgu.n_rows = 2 * (int64_t)F; L.wgu.K = D;
                if ((rc = own(&L.wgu.w, repacked_bytes(t.type, 2 * (int64_t)F, D)))) return rc;
            } else if (L.wgu.type != t.type) return fail(B200_ERR_UNSUPPORTED, "ffn_gate and ffn_up of a layer must share a quant type");
            const size_t grb = raw_row_bytes(t.type, D);
            rc = repack(E, src + (size_t)tr * F * grb, L.wgu.w, t.type, F, F, D, what == "ffn_gate.weight" ? 0 : 1);
            weight_bytes += (uint64_t)F * grb;
        } else return B200_OK;                                                   // a tensor this engine does not use
    } else return B200_OK;
    if (!rc) t_done.insert(t.name);
    return rc;
}

int Model::commit() {
    if (committed) return B200_OK;
    if (t_cursor < t_order.size()) return fail(B200_ERR_STATE, "commit before all tensor bytes were uploaded (" + std::to_string(uploaded) + " of " + std::to_string(meta.data_bytes) + ")");
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    const int D = n_embd, F = n_ff, QD = n_head * head_dim, KVD = n_head_kv * head_dim;
    {   // every tensor the forward pass needs must have arrived
        std::vector<std::string> need = {"token_embd.weight", "output_norm.weight"};
        for (int l = 0; l < n_layer; l++) for (const char* w : {"attn_norm.weight", "attn_q.weight", "attn_k.weight", "attn_v.weight", "attn_output.weight", "ffn_norm.weight", "ffn_gate.weight", "ffn_up.weight", "ffn_down.weight"})
            need.push_back("blk." + std::to_string(l) + "." + w);
        for (auto& nme : need) if (!t_done.count(nme)) return fail(B200_ERR_INVALID, "GGUF: missing tensor " + nme);
    }
    if (!out_w.w) out_w = tok_embd;                                              // tied embeddings

    // rope frequencies (optional rope_freqs.weight = per-frequency divisors, llama-3.1 style)
    std::vector<float> ff(n_rot / 2, 1.0f);
    if (rope_ff.size() == ff.size()) ff = rope_ff;
    std::vector<float> inv(n_rot / 2);
    for (int i = 0; i < n_rot / 2; i++) inv[i] = (float)(std::pow((double)rope_base, -2.0 * i / n_rot) / (double)ff[i]);
    if (int rc = own(&inv_freq, inv.size())) return rc;
    CK(cudaMemcpyAsync(inv_freq, inv.data(), inv.size() * 4, cudaMemcpyHostToDevice, E.stream));
    CK(cudaStreamSynchronize(E.stream));
    release_staging();

    // ---- runtime buffers
    int rc;
    pages_per_seq = (E.max_ctx + 15) / 16;
    // a serving engine (max_seqs > 1) keeps one more sequence slot for the synchronous batch-1 path, so that
    // b200_model_generate / b200_model_eval never write KV pages of a request the scheduler is serving
    sync_slot = E.max_seqs > 1 ? E.max_seqs : 0;
    // KV pool: by default every slot is fully backed (no request can run out of pages); b200_config.kv_pages makes the pool
    // smaller than slots x max_ctx — the scheduler then admits a request only when its pages are free.
    const int table_entries = pages_per_seq * (sync_slot + 1);
    n_pages = E.kv_pages > 0 ? E.kv_pages : table_entries;
    if (n_pages < pages_per_seq) return fail(B200_ERR_INVALID, "kv_pages " + std::to_string(n_pages) + " cannot back one sequence of max_ctx positions (" + std::to_string(pages_per_seq) + " pages)");
    size_t page_elems = (size_t)n_layer * 2 * n_head_kv * 16 * head_dim;
    if ((rc = own(&d_kv, page_elems * n_pages)) || (rc = own(&d_block_table, (size_t)table_entries))) return rc;
    CK(cudaMemset(d_block_table, 0, (size_t)table_entries * 4));       // unbacked entries point at page 0: never dereferenced for a live position
    {
        std::lock_guard<std::mutex> kl(kv_mu);
        kv_host_table.assign(table_entries, 0); kv_slot_pages.assign(sync_slot + 1, {});
        kv_free.clear(); for (int i = 0; i < n_pages; i++) kv_free.push_back(i);      // popped from the back: physical order is the reverse of logical order
    }
    // a tensor-parallel shard runs the same persistent batch-1 kernel on its slice of every matrix; the partial sums of
    // the row-parallel products and the argmax candidates cross NVLink inside the kernel (tp_setup_exchange)
    if (tp) { if (int rc2 = tp_setup_exchange()) return rc2; }
    int max_grp = std::max({n_groups(F), n_groups(QD), 2 * n_groups(D)});
    hown_cap = 2 * ((D / 2 + E.n_sm - 1) / E.n_sm + 2);                       // pairs per CTA slice (worst case: 1 CTA/SM) x 2 rows
    max_segs = 6 * n_layer + 2;                                              // q, k, v, wo, gate/up, down per layer + lm_head
    int need = (head_dim == 128 ? DK_HD128 : DK_HD64) | (tp ? DK_TP : 0) | (getenv("B200_TRACE") ? DK_TRACE : 0);
    for (const Layer& L : layers) for (const Matrix* m : {&L.wq, &L.wk, &L.wv, &L.wo, &L.wgu, &L.wd}) need |= dk_type_bit(m->type);
    need |= dk_type_bit(out_w.type);
    LaunchCfg lc; if (int rc2 = launch_cfg(E, max_grp, hown_cap, max_segs, need, lc)) return rc2;
    grid = lc.grid; ring_bytes = lc.ring_bytes; smem_max_grp = lc.max_grp; smem_bytes = lc.smem; decode_fn = lc.kernel;
    const size_t n_ll16 = (size_t)XREP * ((size_t)D + QD + F) + 2 * (size_t)KVD, n_ll32 = (size_t)QD + (size_t)n_head * 16 * (head_dim + 2) + 2 * (size_t)grid;
    if ((rc = own(&d_ll16, n_ll16)) || (rc = own(&d_ll32, n_ll32)) || (rc = own(&d_logits, (size_t)n_vocab + 2)) || (rc = own(&d_state, (size_t)16)) ||
        (rc = own(&d_out_tokens, (size_t)E.max_ctx + 16)) || (rc = own(&d_forced, (size_t)E.max_ctx + 16)))
        return rc;
    CK(cudaMemset(d_ll16, 0, n_ll16 * 4)); CK(cudaMemset(d_ll32, 0, n_ll32 * 8));     // tag 0 = never valid
    CK(cudaMemset(d_state, 0, 16 * sizeof(int)));

    // ---- decode schedule
    std::vector<Phase> ph;
    auto mk_op = [&](const Matrix& m, int epi, float* out, float scale, int layer, int tp_half = 0) {
        OpDesc o{}; o.tp_half = tp_half; o.w = m.w; o.type = m.type; o.n_pairs = (int)(m.N / 2); o.n_rows = (int)m.n_rows; o.n_grp = n_groups(m.K); o.epi = epi; o.out = out; o.scale = scale; o.layer = layer; return o;
    };
    auto gemv = [&](int K, int pro, const float* src, const float* nw, int layer, std::initializer_list<OpDesc> ops) {
        Phase p{}; p.kind = PH_GEMV; p.K = K; p.pro = pro; p.src = src; p.norm_w = nw; p.layer = layer; p.n_ops = 0; p.total_pairs = 0;
        for (auto& o : ops) { p.op[p.n_ops++] = o; p.total_pairs += o.n_pairs; }
        ph.push_back(p);
    };
    const float rs = res_scale != 0.f ? res_scale : 1.f;
    { Phase p{}; p.kind = PH_EMBED; ph.push_back(p); }
    for (int l = 0; l < n_layer; l++) {
        Layer& L = layers[l];
        gemv(D, PRO_NORM, nullptr, L.attn_norm, l, {mk_op(L.wq, EPI_ROPE_Q, nullptr, 1.f, l), mk_op(L.wk, EPI_ROPE_K, nullptr, 1.f, l), mk_op(L.wv, EPI_V, nullptr, 1.f, l)});
        { Phase p{}; p.kind = PH_ATTN; p.layer = l; ph.push_back(p); }
        gemv(QD, PRO_ATTN, nullptr, nullptr, l, {mk_op(L.wo, EPI_RESID, nullptr, rs, l)});
        gemv(D, PRO_NORM, nullptr, L.ffn_norm, l, {mk_op(L.wgu, EPI_SWIGLU, nullptr, 1.f, l)});
        gemv(F, PRO_ACT, nullptr, nullptr, l, {mk_op(L.wd, EPI_RESID, nullptr, rs, l, 1)});
    }
    gemv(D, PRO_NORM, nullptr, out_norm, 0, {mk_op(out_w, EPI_LOGITS, d_logits, logit_scale != 0.f ? 1.f / logit_scale : 1.f, 0)});
    n_phases = (int)ph.size();
    if ((rc = own(&d_phases, ph.size()))) return rc;
    CK(cudaMemcpy(d_phases, ph.data(), ph.size() * sizeof(Phase), cudaMemcpyHostToDevice));
    committed = true;
    return B200_OK;
}

void Model::fill_params(DecodeParams& P) const {
    P = DecodeParams{};
    P.phases = d_phases; P.n_phases = n_phases; P.sched_layers = n_layer; P.max_segs = max_segs;
    P.n_embd = n_embd; P.n_head = n_head; P.n_head_kv = n_head_kv; P.head_dim = head_dim; P.n_rot = n_rot; P.n_vocab = n_vocab; P.n_vocab_l = n_vocab_l; P.n_splits = n_splits;
    P.rms_eps = rms_eps; P.kq_scale = attn_scale != 0.f ? attn_scale : 1.0f / std::sqrt((float)head_dim); P.emb_scale = emb_scale;
    P.inv_freq = inv_freq; P.embd = tok_embd.w; P.embd_type = tok_embd.type;
    P.logits = d_logits;
    const size_t QD = (size_t)n_head * head_dim, KVD = (size_t)n_head_kv * head_dim;
    P.h_stride = n_embd; P.att_stride = (int)QD; P.act_stride = n_ff;
    P.h_ll = d_ll16; P.att_ll = P.h_ll + (size_t)XREP * n_embd; P.act_ll = P.att_ll + (size_t)XREP * QD; P.kv_ll = P.act_ll + (size_t)XREP * n_ff;
    P.q_ll = d_ll32; P.part_ll = P.q_ll + QD; P.cand_ll = P.part_ll + (size_t)n_head * 16 * (head_dim + 2);
    P.hown_cap = hown_cap; (void)KVD;
    P.n_btab = std::min(pages_per_seq, BTAB_MAX);
    P.kv.pages = d_kv; P.kv.block_table = d_block_table + (size_t)sync_slot * pages_per_seq; P.kv.n_layer = n_layer; P.kv.n_head_kv = n_head_kv; P.kv.head_dim = head_dim; P.kv.pages_per_seq = pages_per_seq; P.kv.n_pages = n_pages;
    P.pos = d_state + 0; P.token = d_state + 1; P.bar_ctr = (unsigned*)(d_state + 2); P.epoch = (unsigned*)(d_state + 3);
    P.out_tokens = d_out_tokens; P.n_steps = 1; P.want_argmax = 1; P.trace = d_trace;
    const Engine& E = engine();
    P.tp_rank = tp ? E.tp_rank : 0; P.tp_world = tp ? E.tp_world : 1; P.vocab_base = tp ? E.tp_rank * n_vocab_l : 0; P.tp_error = tp_error;
    if (tp) for (int r = 0; r < E.tp_world; r++) { P.rx_peer[r] = (uint2*)tp_peer[r]; P.cand_peer[r] = (uint2*)((uint8_t*)tp_peer[r] + tp_rx_bytes); }
}

// Exchange buffers of the tensor-parallel batch-1 kernel: one allocation per rank ([rx: 2 x world x n_embd][candidates:
// world x 2 x grid] tagged 8-byte words, zeroed = never valid), shared with the peers through CUDA IPC handles that
// travel over the NCCL communicator; every rank maps every peer's buffer (NVLink peer stores land in it).
int Model::tp_setup_exchange() {
    Engine& E = engine(); const int W = E.tp_world, me = E.tp_rank;
    if (W > 8) return fail(B200_ERR_UNSUPPORTED, "at most 8 tensor-parallel ranks");
    tp_rx_bytes = (size_t)2 * W * n_embd * sizeof(uint2);
    const size_t bytes = tp_rx_bytes + (size_t)W * 2 * E.n_sm * sizeof(uint2) + 256;
    if (int rc = own(&tp_buf, bytes)) return rc;
    if (int rc = own(&tp_error, (size_t)4)) return rc;
    CK(cudaMemsetAsync(tp_buf, 0, bytes, E.stream)); CK(cudaMemsetAsync(tp_error, 0, 16, E.stream));
    cudaIpcMemHandle_t mine; CK(cudaIpcGetMemHandle(&mine, tp_buf));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    float* d_h = nullptr; if (int rc = dalloc(&d_h, (size_t)(W + 1) * 16)) return rc;
    CK(cudaMemcpyAsync(d_h + (size_t)W * 16, &mine, 64, cudaMemcpyHostToDevice, E.stream));
    if (int rc = E.all_gather(d_h + (size_t)W * 16, d_h, 16)) { cudaFree(d_h); return rc; }      // ordered behind the memsets on every rank
    std::vector<cudaIpcMemHandle_t> all((size_t)W);
    CK(cudaMemcpyAsync(all.data(), d_h, (size_t)W * 64, cudaMemcpyDeviceToHost, E.stream));
    CK(cudaStreamSynchronize(E.stream)); cudaFree(d_h);
    for (int r = 0; r < W; r++) {
        if (r == me) { tp_peer[r] = tp_buf; continue; }
        cudaError_t e = cudaIpcOpenMemHandle(&tp_peer[r], all[(size_t)r], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) return fail(B200_ERR_CUDA, std::string("cudaIpcOpenMemHandle (peer ") + std::to_string(r) + "): " + cudaGetErrorString(e));
    }
    return B200_OK;
}
int Model::tp_check_error() {
    if (!tp || !tp_error) return B200_OK;
    unsigned e = 0; CK(cudaMemcpy(&e, tp_error, 4, cudaMemcpyDeviceToHost));
    if (e) { cudaMemset(tp_error, 0, 4); return fail(B200_ERR_CUDA, "a tensor-parallel peer did not answer inside the decode kernel (ranks must launch the same steps)"); }
    return B200_OK;
}

// Runs `n_steps` decode steps.  forced (host, n_forced ints) are fed first.  reset: start at position 0.
// ------------------------------------------------------------------------------------ KV page allocator
int Model::kv_ensure(int slot, int upto_pos) {
    Engine& E = engine();
    std::lock_guard<std::mutex> kl(kv_mu);
    if (slot < 0 || slot >= (int)kv_slot_pages.size() || upto_pos < 0 || upto_pos >= E.max_ctx) return fail(B200_ERR_INVALID, "kv_ensure: slot or position out of range");
    std::vector<int>& pg = kv_slot_pages[slot];
    const int need = upto_pos / 16 + 1, first_new = (int)pg.size();
    if (need <= first_new) return B200_OK;
    if ((int)kv_free.size() < need - first_new) return fail(B200_ERR_NOMEM, "KV page pool exhausted: " + std::to_string(need - first_new) + " pages wanted, " + std::to_string(kv_free.size()) + " free of " + std::to_string(n_pages));
    while ((int)pg.size() < need) { pg.push_back(kv_free.back()); kv_free.pop_back(); kv_host_table[(size_t)slot * pages_per_seq + pg.size() - 1] = pg.back(); }
    // kv_host_table outlives the copy (it is a member; entries of a slot are only rewritten by a later kv_ensure, which the
    // stream orders behind this one), and pageable-memory copies are staged by the runtime before the call returns
    CK(cudaMemcpyAsync(d_block_table + (size_t)slot * pages_per_seq + first_new, kv_host_table.data() + (size_t)slot * pages_per_seq + first_new, (size_t)(need - first_new) * 4, cudaMemcpyHostToDevice, E.stream));
    return B200_OK;
}
void Model::kv_release(int slot) {
    std::lock_guard<std::mutex> kl(kv_mu);
    if (slot < 0 || slot >= (int)kv_slot_pages.size()) return;
    for (int p : kv_slot_pages[slot]) kv_free.push_back(p);
    kv_slot_pages[slot].clear();
}
int Model::kv_free_count() { std::lock_guard<std::mutex> kl(kv_mu); return (int)kv_free.size(); }

int Model::run_steps(int n_steps, const int32_t* forced, int n_forced, bool reset, float* logits_all_dev) {
    Engine& E = engine();
    if (!committed) return fail(B200_ERR_STATE, "model not committed");
    if (n_steps <= 0) return B200_OK;
    if (reset) { int z = 0; CK(cudaMemcpyAsync(d_state, &z, 4, cudaMemcpyHostToDevice, E.stream)); cur_pos = 0; kv_release(sync_slot); }      // a new sequence: the previous one's pages go back to the pool
    if (cur_pos + n_steps > E.max_ctx) return fail(B200_ERR_INVALID, "context overflow: " + std::to_string(cur_pos + n_steps) + " > max_ctx " + std::to_string(E.max_ctx));
    if (int rc = kv_ensure(sync_slot, cur_pos + n_steps - 1)) return rc;
    if (n_forced > 0) CK(cudaMemcpyAsync(d_forced, forced, (size_t)n_forced * 4, cudaMemcpyHostToDevice, E.stream));
    DecodeParams P; fill_params(P); P.n_steps = n_steps;
    // context splits per head: as many (head, split) items as there are CTAs (each split is one memory round trip);
    // split 0's CTA merges the partials of its head
    P.n_splits = std::max(1, std::min({16, std::max(1, grid / n_head), (cur_pos + n_steps + 15) / 16}));
    if (const char* e = getenv("B200_SPLITS")) P.n_splits = std::max(1, std::min({16, atoi(e), std::max(1, grid / n_head)}));
    P.max_inflight = 4; if (const char* e = getenv("B200_INFLIGHT")) P.max_inflight = std::max(1, atoi(e));
    P.quiet = 1; if (const char* e = getenv("B200_QUIET")) P.quiet = atoi(e);
    DecodeRun R{}; R.forced = d_forced; R.n_forced = n_forced; R.logits_all = logits_all_dev;
    LaunchCfg lc; lc.grid = grid; lc.ring_bytes = ring_bytes; lc.max_grp = smem_max_grp; lc.smem = smem_bytes; lc.kernel = decode_fn;
    if (int rc = launch_decode(E, P, R, lc)) return rc;
    cur_pos += n_steps;
    return B200_OK;
}

int Model::generate(const int32_t* prompt, int n_prompt, int n_gen, const b200_sampling* s, int32_t* out, int* n_out, float* ttft_ms, float* decode_ms,
                    int* finish_reason, const std::chrono::steady_clock::time_point* t_submit) {
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    std::lock_guard<std::mutex> lk(run_mu);
    if (!committed) return fail(B200_ERR_STATE, "model '" + id + "' is not loaded (closed or deleted)");
    if (n_prompt < 1 || n_gen < 1) return fail(B200_ERR_INVALID, "need at least one prompt token and one generated token");
    for (int i = 0; i < n_prompt; i++) if (prompt[i] < 0 || prompt[i] >= n_vocab) return fail(B200_ERR_INVALID, "prompt token out of range");
    const bool greedy = !s || s->temperature <= 0.f;
    if (tp && !greedy) return fail(B200_ERR_UNSUPPORTED, "tensor-parallel shard: the synchronous path is greedy only (each rank holds a slice of the logits)");
    const int stop = s ? s->stop_token : -1; const bool ignore_eos = s && s->ignore_eos;
    const bool has_deadline = s && s->timeout_ms > 0;
    const auto deadline = (t_submit ? *t_submit : std::chrono::steady_clock::now()) + std::chrono::milliseconds(has_deadline ? s->timeout_ms : 0);
    auto expired = [&] { return has_deadline && std::chrono::steady_clock::now() > deadline; };
    bool timed_out = false;
    cudaEvent_t e0, e1, e2; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventCreate(&e2));
    int rc = B200_OK; int produced = 0;
    CK(cudaEventRecord(e0, E.stream));
    if (greedy) {
        // prompt + first token, then the remaining tokens in chunks (EOS checked between chunks).  Long prompts:
        // all but the last token go through the tensor-core prefill (KV pages only), the last one through the decode
        // kernel, which produces the first generated token.
        if (prefill_min > 0 && n_prompt - 1 >= prefill_min) {
            { int z = 0; CK(cudaMemcpyAsync(d_state, &z, 4, cudaMemcpyHostToDevice, E.stream)); cur_pos = 0; kv_release(sync_slot); }
            if ((rc = prefill(prompt, n_prompt - 1))) goto done;
            if ((rc = run_steps(1, prompt + n_prompt - 1, 1, false, nullptr))) goto done;
            CK(cudaMemcpyAsync(d_out_tokens + n_prompt - 1, d_out_tokens, 4, cudaMemcpyDeviceToDevice, E.stream));
        } else if ((rc = run_steps(n_prompt, prompt, n_prompt, true, nullptr))) goto done;
        CK(cudaEventRecord(e1, E.stream));
        CK(cudaMemcpyAsync(out, d_out_tokens + n_prompt - 1, 4, cudaMemcpyDeviceToHost, E.stream));
        CK(cudaStreamSynchronize(E.stream));
        produced = 1;
        while (produced < n_gen) {
            if (!ignore_eos && (out[produced - 1] == eos || out[produced - 1] == stop)) break;
            if (expired()) { timed_out = true; break; }
            int chunk = std::min(n_gen - produced, (ignore_eos && !has_deadline) ? n_gen : 32);
            if ((rc = run_steps(chunk, nullptr, 0, false, nullptr))) goto done;
            CK(cudaMemcpyAsync(out + produced, d_out_tokens, (size_t)chunk * 4, cudaMemcpyDeviceToHost, E.stream));
            CK(cudaStreamSynchronize(E.stream));
            int k = 0;
            for (; k < chunk; k++) { produced++; if (!ignore_eos && (out[produced - 1] == eos || out[produced - 1] == stop)) break; }
        }
    } else {
        // sampled: one step per launch with logits, then the sampler kernel picks the token
        if (prefill_min > 0 && n_prompt - 1 >= prefill_min) {
            { int z = 0; CK(cudaMemcpyAsync(d_state, &z, 4, cudaMemcpyHostToDevice, E.stream)); cur_pos = 0; kv_release(sync_slot); }
            if ((rc = prefill(prompt, n_prompt - 1))) goto done;
            if ((rc = run_steps(1, prompt + n_prompt - 1, 1, false, nullptr))) goto done;
        } else if ((rc = run_steps(n_prompt, prompt, n_prompt, true, nullptr))) goto done;
        for (;;) {
            if ((rc = sample_token(E, *this, *s, (uint64_t)produced))) goto done;
            if (produced == 0) CK(cudaEventRecord(e1, E.stream));
            CK(cudaMemcpyAsync(out + produced, d_state + 1, 4, cudaMemcpyDeviceToHost, E.stream));
            CK(cudaStreamSynchronize(E.stream));
            produced++;
            if (produced >= n_gen || (!ignore_eos && (out[produced - 1] == eos || out[produced - 1] == stop))) break;
            if (expired()) { timed_out = true; break; }
            if ((rc = run_steps(1, nullptr, 0, false, nullptr))) goto done;
        }
    }
    CK(cudaEventRecord(e2, E.stream)); CK(cudaEventSynchronize(e2));
    { float a = 0, b = 0; cudaEventElapsedTime(&a, e0, e1); cudaEventElapsedTime(&b, e1, e2); if (ttft_ms) *ttft_ms = a; if (decode_ms) *decode_ms = b; }
    if (n_out) *n_out = produced;
    if (finish_reason) {
        const bool stopped = produced > 0 && !ignore_eos && (out[produced - 1] == eos || out[produced - 1] == stop);
        *finish_reason = stopped ? B200_FINISH_STOP : timed_out ? B200_FINISH_TIMEOUT : B200_FINISH_LENGTH;
    }
done:
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaEventDestroy(e2);
    return rc;
}

int Model::eval(const int32_t* toks, int n, float* logits_all, float* logits_last) {
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    std::lock_guard<std::mutex> lk(run_mu);
    if (!committed) return fail(B200_ERR_STATE, "model '" + id + "' is not loaded (closed or deleted)");
    if (n < 1) return fail(B200_ERR_INVALID, "empty token list");
    for (int i = 0; i < n; i++) if (toks[i] < 0 || toks[i] >= n_vocab) return fail(B200_ERR_INVALID, "token out of range");
    float* d_all = nullptr;                                        // [n][n_vocab_l]: a tensor-parallel rank produces its vocabulary slice
    if (logits_all) { if (int rc = dalloc(&d_all, (size_t)n * n_vocab_l)) return rc; }
    const size_t col0 = tp ? (size_t)E.tp_rank * n_vocab_l : 0;    // the slice lands in its columns of the caller's [n][n_vocab] array
    int rc;
    if (!logits_all && prefill_min > 0 && n - 1 >= prefill_min) {      // only the last position's logits: batched ingestion of the rest
        int z = 0; CK(cudaMemcpyAsync(d_state, &z, 4, cudaMemcpyHostToDevice, E.stream)); cur_pos = 0; kv_release(sync_slot);
        rc = prefill(toks, n - 1);
        if (!rc) rc = run_steps(1, toks + n - 1, 1, false, nullptr);
    } else rc = run_steps(n, toks, n, true, d_all);
    if (!rc) {
        cudaError_t e = cudaStreamSynchronize(E.stream);
        if (e == cudaSuccess && logits_all) e = cudaMemcpy2D(logits_all + col0, (size_t)n_vocab * 4, d_all, (size_t)n_vocab_l * 4, (size_t)n_vocab_l * 4, n, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess && logits_last) e = cudaMemcpy(logits_last + col0, logits_all ? d_all + (size_t)(n - 1) * n_vocab_l : d_logits, (size_t)n_vocab_l * 4, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess && !rc) rc = tp_check_error();
        if (e != cudaSuccess) rc = fail(B200_ERR_CUDA, cudaGetErrorString(e));
    }
    if (d_all) cudaFree(d_all);
    return rc;
}

// algorithmic bytes one decode step reads at context length ctx:
// every matrix except token_embd (unless tied), one token_embd row, the norm vectors, fp16 K and V.
double Model::bytes_per_token(int ctx) const {
    const GgufTensor* te = meta.tensor("token_embd.weight");
    double w = (double)weight_bytes;
    if (meta.tensor("output.weight") || tp) w -= (double)te->nbytes;      // a shard counts token_embd and its lm_head slice separately
    w += (double)te->nbytes / n_vocab;
    return w + (double)ctx * n_layer * 2 * n_head_kv * head_dim * 2;
}

int Model::bench_decode(int ctx_len, int n_steps, int warmup, float* ms_per_token, double* bytes) {
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    std::lock_guard<std::mutex> lk(run_mu);
    if (!committed) return fail(B200_ERR_STATE, "model '" + id + "' is not loaded (closed or deleted)");
    if (ctx_len < 1 || ctx_len + n_steps + warmup > E.max_ctx) return fail(B200_ERR_INVALID, "bench context does not fit max_ctx");
    // prime the KV cache: positions [0, ctx_len) hold whatever the cache holds (attention cost is
    // data-independent); set the position and a start token, then time n_steps greedy steps.
    int st[2] = {ctx_len, 1 % n_vocab};
    CK(cudaMemcpyAsync(d_state, st, 8, cudaMemcpyHostToDevice, E.stream)); cur_pos = ctx_len;
    if (warmup > 0) { if (int rc = run_steps(warmup, nullptr, 0, false, nullptr)) return rc; }
    CK(cudaMemcpyAsync(d_state, st, 8, cudaMemcpyHostToDevice, E.stream)); cur_pos = ctx_len;
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const char* trace_path = getenv("B200_TRACE");                 // debug: per-phase clock64 stamps of 4 CTAs
    size_t trace_n = (size_t)n_steps * n_phases * grid * 8;
    if (trace_path) { CK(cudaMalloc((void**)&d_trace, trace_n * 8)); CK(cudaMemsetAsync(d_trace, 0, trace_n * 8, E.stream)); }
#ifdef B200_WTRACE
    long long* d_wt = nullptr; const size_t wt_n = (size_t)n_steps * n_phases * 16 * 8;
    if (const char* wp = getenv("B200_WTRACE")) {
        (void)wp; CK(cudaMalloc((void**)&d_wt, wt_n * 8)); CK(cudaMemsetAsync(d_wt, 0, wt_n * 8, E.stream));
        int cta = 0; if (const char* e = getenv("B200_WTRACE_CTA")) cta = atoi(e);
        CK(cudaMemcpyToSymbolAsync(g_wtrace, &d_wt, sizeof(d_wt), 0, cudaMemcpyHostToDevice, E.stream));
        CK(cudaMemcpyToSymbolAsync(g_wtrace_cta, &cta, sizeof(cta), 0, cudaMemcpyHostToDevice, E.stream));
    }
#endif
    CK(cudaEventRecord(e0, E.stream));
    int rc = run_steps(n_steps, nullptr, 0, false, nullptr);
    CK(cudaEventRecord(e1, E.stream)); CK(cudaEventSynchronize(e1));
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1); cudaEventDestroy(e0); cudaEventDestroy(e1);
#ifdef B200_WTRACE
    if (d_wt) {
        std::vector<long long> w(wt_n); cudaMemcpy(w.data(), d_wt, wt_n * 8, cudaMemcpyDeviceToHost);
        if (FILE* f = fopen(getenv("B200_WTRACE"), "wb")) { int hdr[3] = {n_steps, n_phases, 16}; fwrite(hdr, 4, 3, f); fwrite(w.data(), 8, wt_n, f); fclose(f); }
        long long* z = nullptr; cudaMemcpyToSymbol(g_wtrace, &z, sizeof(z)); cudaFree(d_wt);
    }
#endif
    if (trace_path) {
        std::vector<long long> tr(trace_n); cudaMemcpy(tr.data(), d_trace, trace_n * 8, cudaMemcpyDeviceToHost);
        if (FILE* f = fopen(trace_path, "wb")) { int hdr[3] = {n_steps, n_phases, grid}; fwrite(hdr, 4, 3, f); fwrite(tr.data(), 8, trace_n, f); fclose(f); }
        cudaFree(d_trace); d_trace = nullptr;
    }
    if (!rc) rc = tp_check_error();
    if (ms_per_token) *ms_per_token = ms / n_steps;
    if (bytes) { double b = 0; for (int i = 0; i < n_steps; i++) b += bytes_per_token(ctx_len + i + 1); *bytes = b / n_steps; }
    return rc;
}

int sample_token(Engine& E, Model& m, const b200_sampling& s, uint64_t step) {
    sample_kernel<<<1, SAMPLER_THREADS, 0, E.stream>>>(m.d_logits, m.n_vocab, s.top_k, s.top_p <= 0.f ? 1.f : s.top_p, s.temperature, s.seed, step, m.d_state + 1, nullptr);
    CK(cudaGetLastError());
    E.launches++;
    return B200_OK;
}

// ------------------------------------------------------------------------------------ kernel-level entry points
// test entry: one draw of the sampler (K8) over host logits
int kernel_sample(const float* logits, int V, const b200_sampling& s, uint64_t step, int* token) {
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    if (V < 1 || !(s.temperature > 0.f)) return fail(B200_ERR_INVALID, "need V >= 1 and temperature > 0");
    float* d_l = nullptr; int* d_t = nullptr; int rc;
    if ((rc = dalloc(&d_l, (size_t)V)) || (rc = dalloc(&d_t, (size_t)1))) { cudaFree(d_l); cudaFree(d_t); return rc; }
    cudaMemcpyAsync(d_l, logits, (size_t)V * 4, cudaMemcpyHostToDevice, E.stream);
    sample_kernel<<<1, SAMPLER_THREADS, 0, E.stream>>>(d_l, V, s.top_k, s.top_p <= 0.f ? 1.f : s.top_p, s.temperature, s.seed, step, d_t, nullptr); E.launches++;
    cudaError_t e = cudaMemcpyAsync(token, d_t, 4, cudaMemcpyDeviceToHost, E.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(E.stream);
    cudaFree(d_l); cudaFree(d_t);
    return e == cudaSuccess ? B200_OK : fail(B200_ERR_CUDA, cudaGetErrorString(e));
}

int kernel_dequantize(int type, const void* raw, int64_t n_rows, int K, float* out) {
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    if (!is_streamable(type)) return fail(B200_ERR_UNSUPPORTED, "unsupported ggml type");
    if (n_rows < 0 || K <= 0 || K % 256) return fail(B200_ERR_INVALID, "K must be a positive multiple of 256");
    if (n_rows == 0) return B200_OK;
    int64_t N = n_rows + (n_rows & 1);
    size_t rb = raw_row_bytes(type, K);
    uint8_t *d_raw = nullptr, *d_w = nullptr; float* d_out = nullptr; int rc;
    if ((rc = dalloc(&d_raw, (size_t)N * rb)) || (rc = dalloc(&d_w, repacked_bytes(type, N, K))) || (rc = dalloc(&d_out, (size_t)N * K))) { cudaFree(d_raw); cudaFree(d_w); cudaFree(d_out); return rc; }
    cudaMemsetAsync(d_raw, 0, (size_t)N * rb, E.stream);
    cudaMemcpyAsync(d_raw, raw, (size_t)n_rows * rb, cudaMemcpyHostToDevice, E.stream);
    rc = repack(E, d_raw, d_w, type, N, n_rows, K);
    if (!rc) {
        dequant_repacked_kernel<<<std::min<long long>((N * K + 255) / 256, 148 * 16), 256, 0, E.stream>>>(d_w, type, N, K, d_out); E.launches++;
        cudaError_t e = cudaMemcpyAsync(out, d_out, (size_t)n_rows * K * 4, cudaMemcpyDeviceToHost, E.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(E.stream);
        if (e != cudaSuccess) rc = fail(B200_ERR_CUDA, cudaGetErrorString(e));
    }
    cudaFree(d_raw); cudaFree(d_w); cudaFree(d_out);
    return rc;
}

// ------------------------------------------------------------------------------------ prefill (batched prompt ingestion)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr; cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) fn = (EncodeTiledFn)p;
    }
    return fn;
}
// 2-D bf16 tensor [rows][cols] row-major, box = 64 cols x 128 rows, 128-byte swizzle, OOB rows read as zero
static int make_tmap(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, int box_rows = GEMM_BM) {
    EncodeTiledFn f = encode_tiled();
    if (!f) return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t dims[2] = {cols, rows}, strides[1] = {cols * 2};
    cuuint32_t box[2] = {(cuuint32_t)GEMM_BK, (cuuint32_t)box_rows}, estr[2] = {1, 1};
    CUresult r = f(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled failed: " + std::to_string((int)r));
    return B200_OK;
}
// C[M x N] (+)= A[M x K] . B[N x K]^T   (A, B bf16 device, C fp32 device)
// mode 0: C = scale * A.B^T; 1: C += ...; 2 (persistent kernels only, gemm_bf16_fuses_swiglu): C is a bf16 [M x N/2] matrix of silu(gate) * up
static bool gemm_bf16_fuses_swiglu(const Engine& E, int M, int N) {
    static const bool old_kernel = getenv("B200_GEMM_OLD") != nullptr, off = getenv("B200_NO_SWIGLU_FUSE") != nullptr;
    const int m_tiles = (M + GEMM_BM - 1) / GEMM_BM;
    return !old_kernel && !off && N % 32 == 0 && (long long)m_tiles * ((N + 127) / 128) >= E.n_sm / 2;
}
static int gemm_bf16(Engine& E, const __nv_bfloat16* A, const __nv_bfloat16* B, float* C, int M, int N, int K, int ldc, int mode, float scale) {
    if (K % GEMM_BK) return fail(B200_ERR_INVALID, "GEMM K must be a multiple of 64");
    static const bool old_kernel = getenv("B200_GEMM_OLD") != nullptr;
    static bool attr = false;
    if (!attr) {
        CK(cudaFuncSetAttribute(gemm_bf16_tcgen05, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GEMM_SMEM));
        CK(cudaFuncSetAttribute(gemm_bf16_persistent<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PGemm<256>::SMEM));
        CK(cudaFuncSetAttribute(gemm_bf16_persistent<224>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PGemm<224>::SMEM));
        CK(cudaFuncSetAttribute(gemm_bf16_persistent<192>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PGemm<192>::SMEM));
        CK(cudaFuncSetAttribute(gemm_bf16_persistent<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PGemm<128>::SMEM));
        attr = true;
    }
    CUtensorMap ta, tb;
    if (int rc = make_tmap(&ta, A, (uint64_t)M, (uint64_t)K)) return rc;
    const int m_tiles = (M + GEMM_BM - 1) / GEMM_BM;
    // few rows (a test-sized product): the one-tile-per-CTA kernel
    if (old_kernel || (long long)m_tiles * ((N + 127) / 128) < E.n_sm / 2) {
        if (int rc = make_tmap(&tb, B, (uint64_t)N, (uint64_t)K)) return rc;
        dim3 grid((N + GEMM_BN - 1) / GEMM_BN, (M + GEMM_BM - 1) / GEMM_BM);
        gemm_bf16_tcgen05<<<grid, GEMM_THREADS, GEMM_SMEM, E.stream>>>(ta, tb, C, M, N, K, ldc, mode, scale);
    } else {
        // Tile width: a persistent CTA per SM walks ceil(tiles / SMs) tiles of width BN, so the product costs about
        // rounds x (BN + a fixed per-tile part).  Llama-3-8B, 512 rows: QKV (N 6144) 2 x 128 -> 1 x 192; gate/up (N 28672)
        // 4 x 256 -> 4 x 224; wo / down (N 4096) 1 x 128.
        static const int widths[4] = {128, 192, 224, 256};
        int best = 128; long long best_cost = -1;
        for (int w : widths) {
            const long long tiles = (long long)m_tiles * ((N + w - 1) / w), rounds = (tiles + E.n_sm - 1) / E.n_sm, cost = rounds * (w + 24);
            if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = w; }
        }
        if (int rc = make_tmap(&tb, B, (uint64_t)N, (uint64_t)K, best)) return rc;
        const int tiles = m_tiles * ((N + best - 1) / best), grid = std::min(tiles, E.n_sm);
        if (best == 256) gemm_bf16_persistent<256><<<grid, PG_THREADS, PGemm<256>::SMEM, E.stream>>>(ta, tb, C, M, N, K, ldc, mode, scale);
        else if (best == 224) gemm_bf16_persistent<224><<<grid, PG_THREADS, PGemm<224>::SMEM, E.stream>>>(ta, tb, C, M, N, K, ldc, mode, scale);
        else if (best == 192) gemm_bf16_persistent<192><<<grid, PG_THREADS, PGemm<192>::SMEM, E.stream>>>(ta, tb, C, M, N, K, ldc, mode, scale);
        else gemm_bf16_persistent<128><<<grid, PG_THREADS, PGemm<128>::SMEM, E.stream>>>(ta, tb, C, M, N, K, ldc, mode, scale);
    }
    CK(cudaGetLastError());
    E.launches++;
    return B200_OK;
}

// launch with programmatic stream serialization (see pdl_wait / pdl_launch_dependents in prefill.cuh)
template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*k)(KArgs...), dim3 g, dim3 b, size_t smem, cudaStream_t st, Args&&... args) {
    static const bool off = getenv("B200_NO_PDL") != nullptr;
    cudaLaunchConfig_t cfg = {}; cfg.gridDim = g; cfg.blockDim = b; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[0].val.programmaticStreamSerializationAllowed = off ? 0 : 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, k, std::forward<Args>(args)...);
}

// ---- dequant-in-GEMM (qgemm.cuh): C[b][n] (+)= scale * X[b][:] . W[n][:] for up to three matrices that share X (K-permuted fp16, B <= 64 rows)
struct QMat { const Matrix* m; float* C; };
// `sp` (out): what the consuming row kernel needs to complete the tiles stream-K split across CTAs (sp->ws == nullptr
// when there are none); whole = tile-aligned CTA ranges (nothing to complete).  The partials live in one engine-wide
// buffer: the consumer must run before the next qgemm launch (stream order in dense_rows guarantees it).
static int qgemm(Engine& E, const QMat* mats, int n_mat, const __nv_bfloat16* X, int B, int K, int ldc, int mode, float scale, bool whole, QgSplit* sp) {
    if (B < 1 || B > QG_NB || n_mat < 1 || n_mat > 3) return fail(B200_ERR_INVALID, "qgemm: bad batch or matrix count");
    if (K % QG_KSTEP) return fail(B200_ERR_INVALID, "qgemm: K must be a multiple of 256");
    QGemmParams P{}; int steps = 0, tiles = 0; uint32_t slot = 0;
    P.ks = K / QG_KSTEP;
    *sp = QgSplit{};
    for (int i = 0; i < n_mat; i++) {
        const Matrix& m = *mats[i].m;
        if (!m.tl || m.K != K) return fail(B200_ERR_STATE, "qgemm: matrix has no tile layout or K mismatch");
        QSeg& s = P.seg[i]; s.tl = m.tl; s.C = mats[i].C; s.type = m.type; s.n_tiles = (int)((m.N + QG_ROWS - 1) / QG_ROWS); s.n_rows = (int)m.n_rows;
        s.step0 = steps; s.tile0 = tiles; steps += s.n_tiles * P.ks; tiles += s.n_tiles;
        sp->step0[i] = s.step0; sp->col0[i] = (int)(mats[i].C - mats[0].C);
        slot = std::max<uint32_t>(slot, (uint32_t)qg_blob_bytes(m.type));
    }
    P.n_seg = n_mat; P.total_steps = steps; P.n_tiles = tiles; P.B = B; P.ldc = ldc; P.mode = mode; P.scale = scale; P.slot_bytes = slot; P.whole = whole ? 1 : 0;
    const size_t fixed = qg_smem_bytes(0, 0);
    P.n_raw = (int)std::min<size_t>(QG_MAX_RAW, ((size_t)E.smem_optin - fixed) / slot) & ~1;      // even, see QG_A_STEPS
    if (const char* nr = getenv("B200_QG_NRAW")) P.n_raw = std::max(2, std::min(P.n_raw, atoi(nr)) & ~1);
    if (P.n_raw < 2) return fail(B200_ERR_UNSUPPORTED, "qgemm: shared memory too small");
    if (!E.qg_ws) { if (int rc = dalloc(&E.qg_ws, (size_t)E.n_sm * 2 * QG_NB * QG_ROWS)) return rc; }
    P.ws = E.qg_ws;
    if (const char* dbg = getenv("B200_QG_DEBUG")) P.debug = atoi(dbg);
    static bool attr = false;
    if (!attr) { CK(cudaFuncSetAttribute(qgemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, E.smem_optin)); attr = true; }
    // the activation tensor map always spans QG_NB rows of real memory (callers pass a buffer of at least QG_NB x K; rows
    // >= B hold stale finite-or-not values whose accumulator columns are never stored): a box that hangs over the end of a
    // shorter tensor (zero fill) faulted here when it ran next to the weight ring's bulk copies
    CUtensorMap tx; if (int rc = make_tmap(&tx, X, (uint64_t)QG_NB, (uint64_t)K, QG_NB)) return rc;
    const int G = std::min(E.n_sm, whole ? tiles : steps);              // every CTA gets at least one K-step
    // programmatic dependent launch: the CTAs may start their prologue and fill the weight ring while the preceding row
    // kernel is still running; the kernel waits (griddepcontrol.wait) before touching activations or results
    { cudaError_t le = launch_pdl(qgemm_kernel, dim3(G), dim3(QG_THREADS), qg_smem_bytes(slot, P.n_raw), E.stream, tx, P);
      if (le != cudaSuccess) return fail(B200_ERR_CUDA, std::string("qgemm_kernel launch (") + std::to_string(n_mat) + " matrices, type " + std::to_string(P.seg[0].type) + ", K " + std::to_string(K) +
                                         ", tiles " + std::to_string(tiles) + ", B " + std::to_string(B) + ", raw slots " + std::to_string(P.n_raw) + ", whole " + std::to_string(P.whole) + "): " + cudaGetErrorString(le)); }
    CK(cudaGetLastError());
    E.launches++;
    if (!whole) { sp->ws = E.qg_ws; sp->T = steps; sp->G = G; sp->ks = P.ks; sp->n_seg = n_mat; sp->mode = mode; sp->scale = scale * (1.f / QG_PRESCALE); }
    return B200_OK;
}
static int tile_repack(Engine& E, Matrix& m, uint8_t* dst) {
    const long long total = (long long)tile_bytes(m.type, m.N, m.K);
    tile_repack_kernel<<<(int)std::min<long long>((total + 255) / 256, 148LL * 64), 256, 0, E.stream>>>(m.w, m.type, m.N, m.K, dst);
    CK(cudaGetLastError()); E.launches++; m.tl = dst;
    return B200_OK;
}

static inline int nblk(long long n, int per = 256, int cap = 148 * 16) { return (int)std::max<long long>(1, std::min<long long>((n + per - 1) / per, cap)); }

int Model::ensure_prefill_buffers() {
    if (pf_H) return B200_OK;
    Engine& E = engine();
    const int D = n_embd, F = n_ff, QD = n_head * head_dim, KVD = n_head_kv * head_dim, Mc = PREFILL_CHUNK;
    size_t wmax = std::max({(size_t)(QD + 2 * KVD) * D, (size_t)D * QD, (size_t)2 * F * D, (size_t)D * F});
    int rc;
    if ((rc = own(&pf_H, (size_t)Mc * D)) || (rc = own(&pf_X, (size_t)Mc * std::max(D, QD))) || (rc = own(&pf_Y, (size_t)Mc * std::max(QD + 2 * KVD, 2 * F))) ||
        (rc = own(&pf_Q, (size_t)Mc * QD)) || (rc = own(&pf_ACT, (size_t)Mc * F)) || (rc = own(&pf_W, wmax)) || (rc = own(&pf_toks, (size_t)3 * Mc)) ||
        (rc = own(&bt_logits, (size_t)std::max(1, E.max_seqs) * ((n_vocab_l + 3) & ~3) * (E.tp_world + (E.tp_world > 1 ? 1 : 0)))) || (tp && (rc = own(&pf_T, (size_t)Mc * D))) || (rc = own(&bt_next, (size_t)std::max(1, E.max_seqs))) || (rc = own(&bt_samp, (size_t)std::max(1, E.max_seqs))))
        return rc;
    return B200_OK;
}

// bf16 copies of every matrix (batched decode reads them every step; dequantising per step would cost more than the GEMMs)
int Model::ensure_bf16_weights() {
    if (bf16_ready) return B200_OK;
    Engine& E = engine();
    const int D = n_embd, F = n_ff, QD = n_head * head_dim, KVD = n_head_kv * head_dim;
    auto deq = [&](const Matrix& m, __nv_bfloat16* dst) {
        dequant_bf16_kernel<<<nblk(m.n_rows * (long long)m.K / 16, 256, 148 * 32), 256, 0, E.stream>>>(m.w, m.type, m.n_rows, m.K, dst); E.launches++;
    };
    int rc;
    for (int l = 0; l < n_layer; l++) {
        Layer& L = layers[l];
        if ((rc = own(&L.b_qkv, (size_t)(QD + 2 * KVD) * D)) || (rc = own(&L.b_o, (size_t)D * QD)) || (rc = own(&L.b_gu, (size_t)2 * F * D)) || (rc = own(&L.b_d, (size_t)D * F))) return rc;
        deq(L.wq, L.b_qkv); deq(L.wk, L.b_qkv + (size_t)QD * D); deq(L.wv, L.b_qkv + (size_t)(QD + KVD) * D);
        deq(L.wo, L.b_o); deq(L.wgu, L.b_gu); deq(L.wd, L.b_d);
    }
    if ((rc = own(&b_out, (size_t)n_vocab * D))) return rc;
    deq(out_w, b_out);
    CK(cudaStreamSynchronize(E.stream));
    bf16_ready = true;
    return B200_OK;
}

// tile-layout copies of every matrix (second resident copy of the quantised weights; 180 GB of HBM make room for it):
// the batched decode step then reads quantised bytes once per step instead of 2-byte bf16 values
bool Model::use_qgemm(int M) const { return engine().max_seqs > 1 && M <= QG_NB && !getenv("B200_NO_QGEMM"); }

int Model::ensure_tile_weights() {
    if (tile_ready) return B200_OK;
    Engine& E = engine();
    auto mk = [&](Matrix& m) -> int { uint8_t* p = nullptr; if (int rc = own(&p, tile_bytes(m.type, m.N, m.K))) return rc; return tile_repack(E, m, p); };
    int rc;
    for (int l = 0; l < n_layer; l++) {
        Layer& L = layers[l];
        if ((rc = mk(L.wq)) || (rc = mk(L.wk)) || (rc = mk(L.wv)) || (rc = mk(L.wo)) || (rc = mk(L.wgu)) || (rc = mk(L.wd))) return rc;
    }
    if ((rc = mk(out_w))) return rc;
    CK(cudaStreamSynchronize(E.stream));
    tile_ready = true;
    return B200_OK;
}

__global__ void add_rows_kernel(float* __restrict__ H, const float* __restrict__ T, long long n) {
    for (long long i = (blockIdx.x * (long long)blockDim.x + threadIdx.x) * 4; i < n; i += (long long)gridDim.x * blockDim.x * 4) {
        float4 h = *(float4*)(H + i); const float4 t = *(const float4*)(T + i); h.x += t.x; h.y += t.y; h.z += t.z; h.w += t.w; *(float4*)(H + i) = h;
    }
}
// tensor-parallel residual: complete this rank's partial product in pf_T (stream-K parts), sum it over the ranks, add to H
int Model::tp_reduce_add(int M, const QgSplit& sp) {
    Engine& E = engine(); const int D = n_embd;
    if (sp.ws) { qgemm_finish_kernel<<<nblk((long long)M * D), 256, 0, E.stream>>>(sp, pf_T, D, M, D); E.launches++; }
    if (int rc = E.all_reduce_sum(pf_T, (size_t)M * D)) return rc;
    add_rows_kernel<<<nblk((long long)M * D / 4), 256, 0, E.stream>>>(pf_H, pf_T, (long long)M * D); E.launches++;
    return B200_OK;
}

// One pass of the dense path over M activation rows (a prompt chunk of one sequence, or one row per sequence of a
// decode batch).  want_logits: also run the last layer's FFN, the final norm and the lm_head into bt_logits[M][V].
int Model::dense_rows(const int* d_toks, int M, const int* d_pos, const int* d_slot, int pos0, int slot0, bool want_logits, bool q, const std::vector<RowRun>* runs) {
    Engine& E = engine();
    const int D = n_embd, F = n_ff, QD = n_head * head_dim, KVD = n_head_kv * head_dim, W3 = QD + 2 * KVD;
    const float rs = res_scale != 0.f ? res_scale : 1.f, kq = attn_scale != 0.f ? attn_scale : 1.0f / std::sqrt((float)head_dim);
    KvLayout kv; kv.pages = d_kv; kv.block_table = d_block_table; kv.n_layer = n_layer; kv.n_head_kv = n_head_kv; kv.head_dim = head_dim; kv.pages_per_seq = pages_per_seq; kv.n_pages = n_pages;
    RowMap rm{d_pos, d_slot, pos0, slot0};
    // few rows (a decode batch): weights stay quantised and are expanded inside the GEMM; many rows (a prompt chunk):
    // expand to bf16 once per chunk (or use the cache) and run the dense tensor-core GEMM
    // (q is chosen by the caller from the KIND of pass — decode batch or prompt ingestion — never from what ran before
    // or how rows were packed: a request's result does not depend on the load it ran under)
    if (q) { if (int rc0 = ensure_tile_weights()) return rc0; }
    auto deq = [&](const Matrix& m, __nv_bfloat16* dst) {
        dequant_bf16_kernel<<<nblk(m.n_rows * (long long)m.K / 16, 256, 148 * 32), 256, 0, E.stream>>>(m.w, m.type, m.n_rows, m.K, dst); E.launches++;
    };
    const int rms_threads = std::min(1024, std::max(128, (D / 4 + 31) / 32 * 32));     // one float4 per thread where the row allows
    int rc; QgSplit sp{}, none{};                                    // sp: partial tiles the next row kernel must complete
    embed_rows_kernel<<<nblk((long long)M * D), 256, 0, E.stream>>>(tok_embd.w, tok_embd.type, D, d_toks, M, emb_scale, pf_H); E.launches++;
    for (int l = 0; l < n_layer; l++) {
        Layer& L = layers[l];
        const __nv_bfloat16 *wqkv = L.b_qkv, *wo = L.b_o, *wgu = L.b_gu, *wd = L.b_d;
        CK(launch_pdl(rmsnorm_rows_kernel, dim3(M), dim3(rms_threads), 0, E.stream, pf_H, (const float*)L.attn_norm, D, M, rms_eps, pf_X, q, sp)); E.launches++; sp = none;
        if (q) {
            const QMat qkv[3] = {{&L.wq, pf_Y}, {&L.wk, pf_Y + QD}, {&L.wv, pf_Y + QD + KVD}};
            if ((rc = qgemm(E, qkv, 3, pf_X, M, D, W3, 0, 1.f, false, &sp))) return rc;
        } else {
            if (!bf16_ready) { deq(L.wq, pf_W); deq(L.wk, pf_W + (size_t)QD * D); deq(L.wv, pf_W + (size_t)(QD + KVD) * D); wqkv = pf_W; }
            if ((rc = gemm_bf16(E, pf_X, wqkv, pf_Y, M, W3, D, W3, 0, 1.f))) return rc;
        }
        // prompt chunks made of whole runs of one sequence each: causal attention on the tensor cores (attn_tc.cuh)
        const bool tc_off = getenv("B200_NO_ATTN_TC") != nullptr;       // read per call: tests compare the two attention kernels in one process
        const bool tc = !q && runs && !runs->empty() && !tc_off;        // every prompt chunk, whatever it is made of: a request's numerics must not depend on its neighbours
        // batched decode step, head size 128: attn_decode_tc_kernel (operands swapped: positions are the MMA's M; attn_tc.cuh).
        // B = 64, ctx 512: 4.03 ms per step against 4.81 with attn_rows_kernel (B200_NO_ATTN_TC_DEC=1 keeps the row kernel).
        const bool tc_dec = q && !tc_off && getenv("B200_NO_ATTN_TC_DEC") == nullptr && head_dim == 128 && n_head / n_head_kv <= 8 && M <= PREFILL_CHUNK && pages_per_seq <= FD_BTAB;
        CK(launch_pdl(rope_kv_rows_kernel, dim3(nblk((long long)M * W3 / 2)), dim3(256), 0, E.stream, (const float*)pf_Y, M, rm, l, n_head, n_head_kv, head_dim, n_rot, (const float*)inv_freq, kv, pf_Q, sp, tc || tc_dec)); E.launches++; sp = none;
        if (tc_dec) {
            if (int rc2 = ensure_attn_maps()) return rc2;
            const dim3 g(M, n_head_kv); const int G = n_head / n_head_kv;
            CK(launch_pdl(attn_decode_tc_kernel, g, dim3(FD_THREADS), FdCfg::SMEM, E.stream, *(const CUtensorMap*)tmap_qd, *(const CUtensorMap*)tmap_kv, rm, G, l, n_head, n_head_kv, kq, kv, (__half*)pf_X));
            E.launches++;
        } else if (tc) {
            if (int rc2 = ensure_attn_maps()) return rc2;
            FaTiles ft; int nt = 0;
            auto flush = [&]() {
                if (!nt) return;
                const dim3 g(nt, n_head);
                if (head_dim == 128) attn_prefill_tc_kernel<128><<<g, FA_THREADS, FaCfg<128>::SMEM, E.stream>>>(*(const CUtensorMap*)tmap_q, *(const CUtensorMap*)tmap_kv, ft, l, n_head, n_head_kv, kq, kv, pf_X);
                else attn_prefill_tc_kernel<64><<<g, FA_THREADS, FaCfg<64>::SMEM, E.stream>>>(*(const CUtensorMap*)tmap_q, *(const CUtensorMap*)tmap_kv, ft, l, n_head, n_head_kv, kq, kv, pf_X);
                E.launches++; nt = 0;
            };
            for (const RowRun& r : *runs)
                for (int o = 0; o < r.n; o += FA_BM) {
                    ft.t[nt++] = make_int4(r.row0 + o, std::min(FA_BM, r.n - o), r.pos0 + o, r.slot);
                    if (nt == FA_MAX_TILES) flush();
                }
            flush();
            CK(cudaGetLastError());
        } else {
            const int G = n_head / n_head_kv; const dim3 ag(M * n_head_kv);
#define B200_ATTN(HD_, G_) CK(launch_pdl(attn_rows_kernel<HD_, G_>, ag, dim3(ATTN_WARPS * 32), 0, E.stream, (const __nv_bfloat16*)pf_Q, M, rm, l, n_head, n_head_kv, kq, kv, pf_X, q))
            if (head_dim == 128 && G == 4) B200_ATTN(128, 4); else if (head_dim == 128 && G == 8) B200_ATTN(128, 8);
            else if (head_dim == 128 && G == 2) B200_ATTN(128, 2); else if (head_dim == 128 && G == 1) B200_ATTN(128, 1);
            else if (head_dim == 64 && G == 4) B200_ATTN(64, 4); else if (head_dim == 64 && G == 8) B200_ATTN(64, 8);
            else if (head_dim == 64 && G == 2) B200_ATTN(64, 2); else if (head_dim == 64 && G == 1) B200_ATTN(64, 1);
            else if (head_dim == 128) attn_prefill_kernel<4><<<(M * n_head + 3) / 4, 128, 0, E.stream>>>(pf_Q, M, rm, l, n_head, n_head_kv, kq, kv, pf_X, q);
            else attn_prefill_kernel<2><<<(M * n_head + 3) / 4, 128, 0, E.stream>>>(pf_Q, M, rm, l, n_head, n_head_kv, kq, kv, pf_X, q);
#undef B200_ATTN
            E.launches++;
        }
        // residual GEMM: single GPU adds into H (in the epilogue / the next norm's gather); a tensor-parallel shard holds a K
        // slice, so its product is a partial sum: complete it locally, all-reduce it over NVLink, then add it to H
        float* resid = tp ? pf_T : pf_H; const int rmode = tp ? 0 : 1;
        if (q) { const QMat o{&L.wo, resid}; if ((rc = qgemm(E, &o, 1, pf_X, M, QD, D, rmode, rs, false, &sp))) return rc; }
        else {
            if (!bf16_ready) { deq(L.wo, pf_W); wo = pf_W; }
            if ((rc = gemm_bf16(E, pf_X, wo, resid, M, D, QD, D, rmode, rs))) return rc;
        }
        if (tp && (rc = tp_reduce_add(M, sp))) return rc;
        if (tp) sp = none;
        if (l == n_layer - 1 && !want_logits) break;                  // KV fill only: the last layer's FFN feeds nothing
        CK(launch_pdl(rmsnorm_rows_kernel, dim3(M), dim3(rms_threads), 0, E.stream, pf_H, (const float*)L.ffn_norm, D, M, rms_eps, pf_X, q, sp)); E.launches++; sp = none;
        bool fused_act = false;
        if (q) { const QMat gu{&L.wgu, pf_Y}; if ((rc = qgemm(E, &gu, 1, pf_X, M, D, 2 * F, 0, 1.f, false, &sp))) return rc; }
        else {
            if (!bf16_ready) { deq(L.wgu, pf_W); wgu = pf_W; }
            fused_act = gemm_bf16_fuses_swiglu(E, M, 2 * F);           // SwiGLU in the GEMM's epilogue: no fp32 [M x 2F] round trip, one launch less
            if (fused_act) { if ((rc = gemm_bf16(E, pf_X, wgu, (float*)pf_ACT, M, 2 * F, D, F, 2, 1.f))) return rc; }
            else if ((rc = gemm_bf16(E, pf_X, wgu, pf_Y, M, 2 * F, D, 2 * F, 0, 1.f))) return rc;
        }
        if (!fused_act) { CK(launch_pdl(swiglu_rows_kernel, dim3(nblk((long long)M * F / 4)), dim3(256), 0, E.stream, (const float*)pf_Y, M, F, pf_ACT, q, sp)); E.launches++; sp = none; }
        if (q) { const QMat dn{&L.wd, resid}; if ((rc = qgemm(E, &dn, 1, pf_ACT, M, F, D, rmode, rs, false, &sp))) return rc; }
        else {
            if (!bf16_ready) { deq(L.wd, pf_W); wd = pf_W; }
            if ((rc = gemm_bf16(E, pf_ACT, wd, resid, M, D, F, D, rmode, rs))) return rc;
        }
        if (tp && (rc = tp_reduce_add(M, sp))) return rc;
        if (tp) sp = none;
    }
    if (want_logits) {
        // vocab rows are sharded: each rank computes its slice of the logits into part `rank` of bt_logits ([W][M][ldv]),
        // an all-gather makes every part visible everywhere (argmax / sampling then run replicated)
        const int ldv = (n_vocab_l + 3) & ~3; const float ls = logit_scale != 0.f ? 1.f / logit_scale : 1.f;
        float* mine = tp ? bt_logits + (size_t)E.tp_world * M * ldv : bt_logits;       // scratch behind the gathered parts
        CK(launch_pdl(rmsnorm_rows_kernel, dim3(M), dim3(rms_threads), 0, E.stream, pf_H, (const float*)out_norm, D, M, rms_eps, pf_X, q, sp)); E.launches++; sp = none;
        if (q) { const QMat o{&out_w, mine}; if ((rc = qgemm(E, &o, 1, pf_X, M, D, ldv, 0, ls, true, &sp))) return rc; sp = none; }
        else {
            if (!bf16_ready) return fail(B200_ERR_STATE, "logits through the dense path need the bf16 weight cache");
            if ((rc = gemm_bf16(E, pf_X, b_out, mine, M, n_vocab_l, D, ldv, 0, ls))) return rc;
        }
        if (tp && (rc = E.all_gather(mine, bt_logits, (size_t)M * ldv))) return rc;
    }
    CK(cudaGetLastError());
    return B200_OK;
}

// Ingest toks[0..n) of sequence `slot` at positions pos0.. : fills its KV pages (hidden states are discarded).
int Model::prefill_slot(int slot, const int32_t* toks, int n, int pos0) {
    Engine& E = engine();
    if (!committed) return fail(B200_ERR_STATE, "model '" + id + "' is not loaded (closed or deleted)");
    if (n <= 0) return B200_OK;
    if (pos0 + n > E.max_ctx) return fail(B200_ERR_INVALID, "context overflow in prefill");
    if (int rc = kv_ensure(slot, pos0 + n - 1)) return rc;
    if (int rc = ensure_prefill_buffers()) return rc;
    if (n >= 256) maybe_bf16_cache();                                  // long prompts: keep bf16 copies of the matrices when HBM has room (see maybe_bf16_cache)
    for (int c0 = 0; c0 < n; c0 += PREFILL_CHUNK) {
        const int M = std::min(PREFILL_CHUNK, n - c0);
        CK(cudaMemcpyAsync(pf_toks, toks + c0, (size_t)M * 4, cudaMemcpyHostToDevice, E.stream));
        const std::vector<RowRun> run{{0, M, pos0 + c0, slot}};
        if (int rc = dense_rows(pf_toks, M, nullptr, nullptr, pos0 + c0, slot, false, false, &run)) return rc;
    }
    CK(cudaStreamSynchronize(E.stream));                              // toks is caller memory
    return B200_OK;
}

// Ingest M rows that may belong to several sequences (row i: token toks[i] of slot slots[i] at position pos[i]); rows of
// one sequence must appear in position order.  The continuous-batching scheduler packs waiting prompts into such chunks.
int Model::prefill_rows(const int32_t* toks, const int32_t* slots, const int32_t* pos, int M) {
    Engine& E = engine();
    if (!committed) return fail(B200_ERR_STATE, "model '" + id + "' is not loaded (closed or deleted)");
    if (M < 1 || M > PREFILL_CHUNK) return fail(B200_ERR_INVALID, "prefill chunk size out of range");
    for (int i = 0; i < M; i++) if (toks[i] < 0 || toks[i] >= n_vocab || slots[i] < 0 || slots[i] >= E.max_seqs || pos[i] < 0 || pos[i] >= E.max_ctx) return fail(B200_ERR_INVALID, "bad prefill row");
    for (int i = 0; i < M; i++) if (int rc = kv_ensure(slots[i], pos[i])) return rc;
    if (int rc = ensure_prefill_buffers()) return rc;
    if (E.max_seqs > 1) maybe_bf16_cache();
    const int Mc = PREFILL_CHUNK;
    CK(cudaMemcpyAsync(pf_toks, toks, (size_t)M * 4, cudaMemcpyHostToDevice, E.stream));
    CK(cudaMemcpyAsync(pf_toks + Mc, pos, (size_t)M * 4, cudaMemcpyHostToDevice, E.stream));
    CK(cudaMemcpyAsync(pf_toks + 2 * Mc, slots, (size_t)M * 4, cudaMemcpyHostToDevice, E.stream));
    std::vector<RowRun> runs;                                          // maximal runs of consecutive positions of one sequence
    for (int i = 0; i < M; i++) {
        if (!runs.empty() && runs.back().slot == slots[i] && runs.back().pos0 + runs.back().n == pos[i]) runs.back().n++;
        else runs.push_back(RowRun{i, 1, pos[i], slots[i]});
    }
    if (int rc = dense_rows(pf_toks, M, pf_toks + Mc, pf_toks + 2 * Mc, 0, 0, false, false, &runs)) return rc;
    CK(cudaStreamSynchronize(E.stream));
    return B200_OK;
}

// tensor maps of the tensor-core prompt attention: the query buffer (fp16 view of pf_Q) and the whole paged KV cache as
// one 2-D fp16 tensor [(layer, page, K|V, kv head, position-in-page)][head_dim], boxes of one page x 64 columns
int Model::ensure_attn_maps() {
    if (tmap_q) return B200_OK;
    EncodeTiledFn f = encode_tiled();
    if (!f) return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    static_assert(sizeof(CUtensorMap) == 128, "tensor map size");
    tmap_storage.assign(3 * sizeof(CUtensorMap) + 64, 0);
    uint8_t* al = (uint8_t*)(((uintptr_t)tmap_storage.data() + 63) & ~(uintptr_t)63);
    CUtensorMap* mq = (CUtensorMap*)al; CUtensorMap* mkv = mq + 1; CUtensorMap* mqd = mq + 2;
    const uint64_t QD = (uint64_t)n_head * head_dim;
    {
        cuuint64_t dims[2] = {QD, (cuuint64_t)PREFILL_CHUNK}, strides[1] = {QD * 2}; cuuint32_t box[2] = {64, (cuuint32_t)FA_BM}, estr[2] = {1, 1};
        CUresult r = f(mq, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, pf_Q, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled (Q) failed: " + std::to_string((int)r));
    }
    {
        const uint64_t rows = (uint64_t)n_layer * n_pages * 2 * n_head_kv * 16;
        cuuint64_t dims[2] = {(cuuint64_t)head_dim, rows}, strides[1] = {(cuuint64_t)head_dim * 2}; cuuint32_t box[2] = {64, 16}, estr[2] = {1, 1};
        CUresult r = f(mkv, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d_kv, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled (KV) failed: " + std::to_string((int)r));
    }
    {   // Q of a decode step seen as [token * head][head_dim]: the G heads of a kv group are G consecutive rows
        cuuint64_t dims[2] = {(cuuint64_t)head_dim, (cuuint64_t)PREFILL_CHUNK * n_head}, strides[1] = {(cuuint64_t)head_dim * 2}; cuuint32_t box[2] = {64, (cuuint32_t)FD_N}, estr[2] = {1, 1};      // the G heads of a kv group, padded to the MMA's N = 16
        CUresult r = f(mqd, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, pf_Q, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(B200_ERR_CUDA, "cuTensorMapEncodeTiled (Q, decode view) failed: " + std::to_string((int)r));
    }
    CK(cudaFuncSetAttribute(attn_prefill_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FaCfg<128>::SMEM));
    CK(cudaFuncSetAttribute(attn_prefill_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FaCfg<64>::SMEM));
    CK(cudaFuncSetAttribute(attn_decode_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FdCfg::SMEM));
    tmap_q = mq; tmap_kv = mkv; tmap_qd = mqd;
    return B200_OK;
}

// A serving engine (max_seqs > 1) ingests prompts all the time: keep bf16 copies of the matrices when HBM has room
// (8B: 16 GB) instead of expanding them for every chunk.  Tried once; without room the per-chunk expansion stays.
void Model::maybe_bf16_cache() {
    if (bf16_ready || bf16_tried) return;
    bf16_tried = true;
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) return;
    const size_t D = n_embd, F = n_ff, QD = (size_t)n_head * head_dim, KVD = (size_t)n_head_kv * head_dim;
    const size_t need = 2 * ((size_t)n_layer * ((QD + 2 * KVD) * D + D * QD + 2 * F * D + D * F) + (size_t)n_vocab * D);
    if (free_b > need + ((size_t)8 << 30)) { if (ensure_bf16_weights() != B200_OK) cudaGetLastError(); }
}

int Model::prefill(const int32_t* toks, int n) {                      // batch-1 path: sequence slot 0, continues at cur_pos
    Engine& E = engine();
    if (int rc = prefill_slot(sync_slot, toks, n, cur_pos)) return rc;
    cur_pos += n;
    int st = cur_pos; CK(cudaMemcpyAsync(d_state, &st, 4, cudaMemcpyHostToDevice, E.stream));
    CK(cudaStreamSynchronize(E.stream));
    return B200_OK;
}

// One decode step for B sequences: row i feeds token toks[i] of sequence slot slots[i] at position pos[i];
// next[i] = greedy next token.  Rows are independent (a row's result does not depend on which other rows are present).
int Model::batch_step(const int32_t* toks, const int32_t* slots, const int32_t* pos, int B, int32_t* next, float* logits, const b200_sampling* samp, const int32_t* draw) {
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    if (!committed) return fail(B200_ERR_STATE, "model '" + id + "' is not loaded (closed or deleted)");
    if (B < 1 || B > E.max_seqs || B > PREFILL_CHUNK) return fail(B200_ERR_INVALID, "batch size out of range");
    for (int i = 0; i < B; i++) if (toks[i] < 0 || toks[i] >= n_vocab || slots[i] < 0 || slots[i] >= E.max_seqs || pos[i] < 0 || pos[i] >= E.max_ctx) return fail(B200_ERR_INVALID, "bad batch row");
    for (int i = 0; i < B; i++) if (int rc = kv_ensure(slots[i], pos[i])) return rc;
    if (int rc = ensure_prefill_buffers()) return rc;
    if (!use_qgemm(B)) { if (int rc = ensure_bf16_weights()) return rc; }
    const int Mc = PREFILL_CHUNK;
    CK(cudaMemcpyAsync(pf_toks, toks, (size_t)B * 4, cudaMemcpyHostToDevice, E.stream));
    CK(cudaMemcpyAsync(pf_toks + Mc, pos, (size_t)B * 4, cudaMemcpyHostToDevice, E.stream));
    CK(cudaMemcpyAsync(pf_toks + 2 * Mc, slots, (size_t)B * 4, cudaMemcpyHostToDevice, E.stream));
    const int ldv = (n_vocab_l + 3) & ~3, Wp = E.tp_world;
    {   // per-row sampling requests (temperature 0 = keep the argmax); the draw counter is the number of tokens the
        // request has generated so far, the same counter the batch-1 path uses, so a seeded request draws the same
        // tokens whether it runs alone or inside a batch
        std::vector<RowSampling> rs((size_t)B);
        for (int i = 0; i < B; i++) {
            RowSampling r{}; if (samp && samp[i].temperature > 0.f) { r.temperature = samp[i].temperature; r.top_p = samp[i].top_p; r.top_k = samp[i].top_k; r.seed = samp[i].seed; r.step = draw ? (uint64_t)draw[i] : 0; }
            rs[i] = r;
        }
        CK(cudaMemcpyAsync(bt_samp, rs.data(), (size_t)B * sizeof(RowSampling), cudaMemcpyHostToDevice, E.stream));   // pageable source: staged before the call returns
    }
    if (int rc = step_enqueue(B)) return rc;
    CK(cudaMemcpyAsync(next, bt_next, (size_t)B * 4, cudaMemcpyDeviceToHost, E.stream));
    if (logits) for (int p = 0; p < Wp; p++)
        CK(cudaMemcpy2DAsync(logits + (size_t)p * n_vocab_l, (size_t)n_vocab * 4, bt_logits + (size_t)p * B * ldv, (size_t)ldv * 4, (size_t)n_vocab_l * 4, B, cudaMemcpyDeviceToHost, E.stream));
    CK(cudaStreamSynchronize(E.stream));
    return B200_OK;
}

// Enqueues one batched decode step over B rows whose tokens / positions / slots / sampling requests are already in
// the device buffers (pf_toks, bt_samp): direct launches the first time a batch size is seen, a CUDA graph afterwards.
int Model::step_enqueue(int B) {
    Engine& E = engine();
    const int Mc = PREFILL_CHUNK, ldv = (n_vocab_l + 3) & ~3, Wp = E.tp_world;
    const bool q = use_qgemm(B);
    auto body = [&]() -> int {
        if (int rc = dense_rows(pf_toks, B, pf_toks + Mc, pf_toks + 2 * Mc, 0, 0, true, q)) return rc;
        CK(launch_pdl(argmax_rows_kernel, dim3(B), dim3(1024), 0, E.stream, (const float*)bt_logits, n_vocab_l, ldv, Wp, (size_t)B * ldv, bt_next)); E.launches++;
        sample_rows_kernel<<<B, SAMPLER_THREADS, 0, E.stream>>>((const float*)bt_logits, n_vocab, n_vocab_l, ldv, (size_t)B * ldv, (const RowSampling*)bt_samp, bt_next); E.launches++;
        CK(cudaGetLastError());
        return B200_OK;
    };
    // The step is a fixed chain of ~9 launches per layer whose arguments depend only on B (tokens, positions and slots are
    // read from device buffers): the second time a batch size is seen the chain is captured into a CUDA graph and replayed
    // from then on.  With the weights split over 8 GPUs the kernels are short enough that issuing ~1400 launches (NCCL
    // included) from the host was the bound of the step.
    static const bool graphs_on = getenv("B200_NO_GRAPH") == nullptr;
    if (step_graph.empty()) { step_graph.assign(QG_NB + 1, nullptr); step_seen.assign(QG_NB + 1, 0); step_launches.assign(QG_NB + 1, 0); }
    if (q && graphs_on && !graph_failed && step_graph[B]) {
        CK(cudaGraphLaunch(step_graph[B], E.stream)); E.launches += step_launches[B];
    } else if (q && graphs_on && !graph_failed && step_seen[B]) {
        cudaGraph_t g = nullptr; const uint64_t l0 = E.launches;
        cudaError_t ce = cudaStreamBeginCapture(E.stream, cudaStreamCaptureModeThreadLocal);
        int rc = ce == cudaSuccess ? body() : B200_ERR_CUDA;
        cudaError_t ee = cudaStreamEndCapture(E.stream, &g);
        if (rc == B200_OK && ee == cudaSuccess && g && cudaGraphInstantiate(&step_graph[B], g, 0) == cudaSuccess) {
            step_launches[B] = (int)(E.launches - l0);
            CK(cudaGraphLaunch(step_graph[B], E.stream));
        } else {                                                   // capture not possible here: stay with direct launches
            graph_failed = true; step_graph[B] = nullptr; cudaGetLastError(); E.launches = l0;
            if (int rc2 = body()) return rc2;
        }
        if (g) cudaGraphDestroy(g);
    } else {
        if (int rc = body()) return rc;
        step_seen[B] = 1;                                          // buffers, tile weights and function attributes now exist
    }
    return B200_OK;
}

// Device-timed batched step (bench.py "batched" roofline): B sequences at context ctx, `iters` steps between two CUDA
// events on the engine stream, nothing crossing PCIe inside the timed region.  The KV pages hold whatever they hold
// (the cost of attention does not depend on the values), every step re-runs position ctx.
// bytes = algorithmic HBM bytes of one step: every matrix except token_embd once (the batch shares the weights), B
// embedding rows, the norm vectors, and fp16 K/V of ctx+1 positions for each of the B sequences.
int Model::bench_batch(int B, int ctx, int iters, int warmup, float* ms_per_step, double* bytes) {
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    std::lock_guard<std::mutex> lk(run_mu);
    if (!committed) return fail(B200_ERR_STATE, "model '" + id + "' is not loaded (closed or deleted)");
    if (B < 1 || B > E.max_seqs || B > QG_NB || ctx < 1 || ctx + 1 > E.max_ctx || iters < 1) return fail(B200_ERR_INVALID, "bench_batch: batch, context or iteration count out of range");
    if (int rc = ensure_prefill_buffers()) return rc;
    if (!use_qgemm(B)) { if (int rc = ensure_bf16_weights()) return rc; }
    const int Mc = PREFILL_CHUNK;
    std::vector<int32_t> toks(B), pos(B, ctx), slots(B);
    for (int i = 0; i < B; i++) { toks[i] = (int)((1000003u * (unsigned)(i + 1)) % (unsigned)n_vocab); slots[i] = i; if (int rc = kv_ensure(i, ctx)) return rc; }
    std::vector<RowSampling> rs((size_t)B, RowSampling{});
    CK(cudaMemcpyAsync(pf_toks, toks.data(), (size_t)B * 4, cudaMemcpyHostToDevice, E.stream));
    CK(cudaMemcpyAsync(pf_toks + Mc, pos.data(), (size_t)B * 4, cudaMemcpyHostToDevice, E.stream));
    CK(cudaMemcpyAsync(pf_toks + 2 * Mc, slots.data(), (size_t)B * 4, cudaMemcpyHostToDevice, E.stream));
    CK(cudaMemcpyAsync(bt_samp, rs.data(), (size_t)B * sizeof(RowSampling), cudaMemcpyHostToDevice, E.stream));
    for (int i = 0; i < std::max(warmup, 2); i++) if (int rc = step_enqueue(B)) return rc;      // >= 2: the second pass builds the graph
    CK(cudaStreamSynchronize(E.stream));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, E.stream));
    int rc = B200_OK;
    for (int i = 0; i < iters && !rc; i++) rc = step_enqueue(B);
    CK(cudaEventRecord(e1, E.stream)); CK(cudaEventSynchronize(e1));
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1); cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (ms_per_step) *ms_per_step = ms / iters;
    if (bytes) {
        const GgufTensor* te = meta.tensor("token_embd.weight");
        double w = (double)weight_bytes;
        if (meta.tensor("output.weight") || tp) w -= (double)te->nbytes;          // a shard counts token_embd and its output slice separately
        *bytes = w + (double)B * te->nbytes / n_vocab + (double)B * (ctx + 1) * n_layer * 2 * n_head_kv * head_dim * 2;
    }
    return rc;
}

// test entry: y[b][n] = sum_k x[b][k] * W[n][k] through the dequant-in-GEMM kernel (raw ggml rows from the host)
int kernel_qgemm(int type, const void* raw, int64_t N, int K, const float* x, int B, float* y, int iters, float* ms_out) {
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    if (!is_streamable(type)) return fail(B200_ERR_UNSUPPORTED, "unsupported ggml type");
    if (N <= 0 || K <= 0 || K % 256 || B < 1 || B > QG_NB) return fail(B200_ERR_INVALID, "K must be a multiple of 256 and 1 <= B <= 64");
    const int64_t Np = N + (N & 1); const size_t rb = raw_row_bytes(type, K);
    Matrix m; m.type = type; m.N = Np; m.n_rows = N; m.K = K;
    uint8_t *d_raw = nullptr, *d_tl = nullptr; __nv_bfloat16* d_x = nullptr; float* d_y = nullptr; int rc;
    auto cleanup = [&]() { cudaFree(d_raw); cudaFree(m.w); cudaFree(d_tl); cudaFree(d_x); cudaFree(d_y); };
    if ((rc = dalloc(&d_raw, (size_t)N * rb)) || (rc = dalloc(&m.w, repacked_bytes(type, Np, K))) || (rc = dalloc(&d_tl, tile_bytes(type, Np, K))) ||
        (rc = dalloc(&d_x, (size_t)QG_NB * K)) || (rc = dalloc(&d_y, (size_t)B * N))) { cleanup(); return rc; }
    cudaMemsetAsync(d_x, 0, (size_t)QG_NB * K * 2, E.stream);
    std::vector<__nv_bfloat16> hx((size_t)B * K);
    for (int b = 0; b < B; b++) for (int k = 0; k < K; k++) { const __half h = __float2half_rn(x[(size_t)b * K + kperm4(k)]); memcpy(&hx[(size_t)b * K + k], &h, 2); }
    cudaMemcpyAsync(d_raw, raw, (size_t)N * rb, cudaMemcpyHostToDevice, E.stream);
    cudaMemcpyAsync(d_x, hx.data(), hx.size() * 2, cudaMemcpyHostToDevice, E.stream);
    cudaMemsetAsync(d_y, 0, (size_t)B * N * 4, E.stream);
    if (!(rc = repack(E, d_raw, m.w, type, Np, N, K)) && !(rc = tile_repack(E, m, d_tl))) {
        const QMat qm{&m, d_y}; QgSplit sp{};
        const bool whole = getenv("B200_QG_WHOLE") != nullptr;
        rc = qgemm(E, &qm, 1, d_x, B, K, (int)N, 0, 1.f, whole, &sp);
        if (!rc && sp.ws) { qgemm_finish_kernel<<<nblk((long long)B * N), 256, 0, E.stream>>>(sp, d_y, (int)N, B, (int)N); E.launches++; }
        if (!rc && iters > 1) {                                    // timing: repeat the launch pair, mean per iteration
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            cudaEventRecord(e0, E.stream);
            for (int i = 0; i < iters && !rc; i++) rc = qgemm(E, &qm, 1, d_x, B, K, (int)N, 0, 1.f, whole, &sp);
            cudaEventRecord(e1, E.stream); cudaEventSynchronize(e1);
#ifdef B200_QG_TRACE
            if (const char* tp_ = getenv("B200_QG_TRACE")) {          // the last launch's stamps of CTA 0
                std::vector<long long> tr(6 * 64 * 8); cudaMemcpyFromSymbol(tr.data(), g_qgt, tr.size() * 8);
                if (FILE* f = fopen(tp_, "wb")) { fwrite(tr.data(), 8, tr.size(), f); fclose(f); }
            }
#endif
            float ms = 0.f; cudaEventElapsedTime(&ms, e0, e1); if (ms_out) *ms_out = ms / iters;
            cudaEventDestroy(e0); cudaEventDestroy(e1);
        }
    }
    if (!rc) {
        cudaError_t e = cudaMemcpyAsync(y, d_y, (size_t)B * N * 4, cudaMemcpyDeviceToHost, E.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(E.stream);
        if (e != cudaSuccess) rc = fail(B200_ERR_CUDA, cudaGetErrorString(e));
    }
    cudaStreamSynchronize(E.stream);
    cleanup();
    return rc;
}

// test entry: C = A . B^T with host fp32 inputs rounded to bf16 on the way in
int kernel_gemm_bf16(const float* A, const float* B, float* C, int M, int N, int K) {
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    if (M <= 0 || N <= 0 || K <= 0 || K % GEMM_BK) return fail(B200_ERR_INVALID, "K must be a multiple of 64");
    std::vector<__nv_bfloat16> ha((size_t)M * K), hb((size_t)N * K);
    for (size_t i = 0; i < ha.size(); i++) ha[i] = __float2bfloat16(A[i]);
    for (size_t i = 0; i < hb.size(); i++) hb[i] = __float2bfloat16(B[i]);
    __nv_bfloat16 *dA = nullptr, *dB = nullptr; float* dC = nullptr; int rc;
    if ((rc = dalloc(&dA, ha.size())) || (rc = dalloc(&dB, hb.size())) || (rc = dalloc(&dC, (size_t)M * N))) { cudaFree(dA); cudaFree(dB); cudaFree(dC); return rc; }
    cudaMemcpyAsync(dA, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice, E.stream);
    cudaMemcpyAsync(dB, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice, E.stream);
    cudaMemsetAsync(dC, 0, (size_t)M * N * 4, E.stream);
    rc = gemm_bf16(E, dA, dB, dC, M, N, K, N, 0, 1.f);
    if (!rc) {
        cudaError_t e = cudaMemcpyAsync(C, dC, (size_t)M * N * 4, cudaMemcpyDeviceToHost, E.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(E.stream);
        if (e != cudaSuccess) rc = fail(B200_ERR_CUDA, cudaGetErrorString(e));
    }
    cudaFree(dA); cudaFree(dB); cudaFree(dC);
    return rc;
}

// ------------------------------------------------------------------------------------ allocation-free dequantise
// Same two kernels as kernel_dequantize (repack into the engine layout, dequantise back), but the intermediate
// engine-layout copy lives in a fixed array inside the module image instead of a cudaMalloc'd buffer, and the
// caller supplies device pointers for the raw blocks and the fp32 result.  Nothing is allocated or freed here.
// The array is one per GPU that loads this library and stays resident until the process exits.
static __device__ __align__(256) unsigned char g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];
static unsigned char* g_noalloc_ptr[64];          // device address of the array, looked up once per device
static std::mutex g_noalloc_mu;

int kernel_dequantize_noalloc(int type, const void* d_raw, int64_t n_rows, int K, float* d_out) {
    Engine& E = engine();
    if (!E.ready) return fail(B200_ERR_STATE, "b200_init must run before b200_dequantize_noalloc (this entry point creates nothing, not even streams)");
    if (!is_streamable(type)) return fail(B200_ERR_UNSUPPORTED, "unsupported ggml type");
    if (n_rows < 0 || K <= 0 || K % 256) return fail(B200_ERR_INVALID, "K must be a positive multiple of 256");
    if (n_rows == 0) return B200_OK;
    if (!d_raw || !d_out) return fail(B200_ERR_INVALID, "null device pointer");
    const int64_t N = n_rows + (n_rows & 1);
    const size_t need = repacked_bytes(type, N, K);
    if (need > (size_t)B200_NOALLOC_SCRATCH_BYTES)
        return fail(B200_ERR_INVALID, "b200_dequantize_noalloc: " + std::to_string(need) + " bytes of scratch needed, the built-in buffer holds " +
                                          std::to_string((size_t)B200_NOALLOC_SCRATCH_BYTES) + " (use b200_dequantize, which allocates)");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);                    // host callers of this library take turns on the shared array
    if (E.device < 0 || E.device >= 64) return fail(B200_ERR_INVALID, "device ordinal out of range");
    CK(cudaSetDevice(E.device));
    if (!g_noalloc_ptr[E.device]) {
        // First use on this device: resolve the array's device address.  In host code the array's name is only a
        // handle for the runtime, not something a kernel can dereference.
        unsigned char* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(B200_ERR_CUDA, std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e));
        g_noalloc_ptr[E.device] = p;
    }
    unsigned char* scratch = g_noalloc_ptr[E.device];
    // Nothing in the array has to start at zero: repack_kernel writes every byte of the `need` bytes (padding rows
    // and padding groups included) before dequant_repacked_kernel reads them, so leftovers of earlier calls are inert.
    // Queued on the engine's utility stream (created by b200_init), never on the generation stream: the scheduler may be
    // capturing that one into a CUDA graph at this moment, and a foreign launch or a synchronise would break the capture.
    if (int rc = repack(E, (const uint8_t*)d_raw, scratch, type, N, n_rows, K, -1, 0, E.aux_stream)) return rc;
    dequant_repacked_kernel<<<(int)std::min<long long>((n_rows * (long long)K + 255) / 256, 148 * 16), 256, 0, E.aux_stream>>>(scratch, type, n_rows, K, d_out);
    CK(cudaGetLastError());
    E.launches++;
    CK(cudaStreamSynchronize(E.aux_stream));                         // the array is free again when this returns
    return B200_OK;
}

int kernel_gemv(int type, const void* raw, int64_t N, int K, const float* x, float* y, int iters, float* ms_out) {
    Engine& E = engine(); if (int rc = E.ensure()) return rc;
    if (!is_streamable(type)) return fail(B200_ERR_UNSUPPORTED, "unsupported ggml type");
    if (N <= 0 || N % 2 || K <= 0 || K % 256) return fail(B200_ERR_INVALID, "N must be even and K a multiple of 256");
    if (iters < 1) iters = 1;
    size_t rb = raw_row_bytes(type, K);
    uint8_t *d_raw = nullptr, *d_w = nullptr; float *d_x = nullptr, *d_y = nullptr; uint2* d_cv = nullptr; int *d_st = nullptr, *d_ot = nullptr; Phase* d_ph = nullptr;
    int rc = B200_OK;
    LaunchCfg lc;
    auto cleanup = [&]() { cudaFree(d_raw); cudaFree(d_w); cudaFree(d_x); cudaFree(d_y); cudaFree(d_cv); cudaFree(d_st); cudaFree(d_ot); cudaFree(d_ph); };
    if ((rc = launch_cfg(E, n_groups(K), 4, 4, DK_HD128 | dk_type_bit(type), lc))) return rc;
    if ((rc = dalloc(&d_raw, (size_t)N * rb)) || (rc = dalloc(&d_w, repacked_bytes(type, N, K))) || (rc = dalloc(&d_x, (size_t)K)) || (rc = dalloc(&d_y, (size_t)N)) ||
        (rc = dalloc(&d_cv, (size_t)2 * lc.grid)) || (rc = dalloc(&d_st, (size_t)16)) || (rc = dalloc(&d_ot, (size_t)iters + 1)) || (rc = dalloc(&d_ph, (size_t)1))) { cleanup(); return rc; }
    cudaMemcpyAsync(d_raw, raw, (size_t)N * rb, cudaMemcpyHostToDevice, E.stream);
    cudaMemcpyAsync(d_x, x, (size_t)K * 4, cudaMemcpyHostToDevice, E.stream);
    cudaMemsetAsync(d_st, 0, 64, E.stream);
    rc = repack(E, d_raw, d_w, type, N, N, K);
    if (!rc) {
        Phase p{}; p.kind = PH_GEMV; p.K = K; p.pro = PRO_PLAIN; p.src = d_x; p.n_ops = 1; p.total_pairs = (int)(N / 2);
        p.op[0].w = d_w; p.op[0].type = type; p.op[0].n_pairs = (int)(N / 2); p.op[0].n_rows = (int)N; p.op[0].n_grp = n_groups(K); p.op[0].epi = EPI_STORE; p.op[0].out = d_y; p.op[0].scale = 1.f;
        cudaMemcpyAsync(d_ph, &p, sizeof p, cudaMemcpyHostToDevice, E.stream);
        DecodeParams P{}; P.phases = d_ph; P.n_phases = 1; P.sched_layers = 0; P.max_segs = 4; P.n_embd = K; P.n_head = 1; P.n_head_kv = 1; P.head_dim = 128; P.n_vocab = (int)N; P.n_splits = 1;
        P.cand_ll = d_cv; P.hown_cap = 4; P.pos = d_st; P.token = d_st + 1; P.bar_ctr = (unsigned*)(d_st + 2); P.epoch = (unsigned*)(d_st + 3); P.out_tokens = d_ot;
        P.logits = d_y; P.n_steps = 1; P.max_inflight = 4; if (const char* e = getenv("B200_INFLIGHT")) P.max_inflight = std::max(1, atoi(e));
        // every step "forced" (token 0): no argmax traffic; forced[] = d_ot, zeroed, iters+1 ints
        cudaMemsetAsync(d_ot, 0, ((size_t)iters + 1) * 4, E.stream);
        DecodeRun R{}; R.forced = d_ot; R.n_forced = iters + 1;
        rc = launch_decode(E, P, R, lc);                                   // warm-up + correctness
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        if (!rc) {
            P.n_steps = iters;
            cudaEventRecord(e0, E.stream);
            rc = launch_decode(E, P, R, lc);
            cudaEventRecord(e1, E.stream);
        }
        cudaError_t e = cudaMemcpyAsync(y, d_y, (size_t)N * 4, cudaMemcpyDeviceToHost, E.stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(E.stream);
        if (e != cudaSuccess && !rc) rc = fail(B200_ERR_CUDA, cudaGetErrorString(e));
        float ms = 0; if (!rc) cudaEventElapsedTime(&ms, e0, e1);
        if (ms_out) *ms_out = ms / iters;
        cudaEventDestroy(e0); cudaEventDestroy(e1);
    }
    cleanup();
    return rc;
}

int layout_selftest() {
    {   // K-quant scale re-encoding is an exact bijection on the 96 bits
        uint32_t st = 12345u;
        for (int it = 0; it < 4096; it++) {
            uint8_t q[12], e[12], r[12];
            for (int i = 0; i < 12; i++) { st = st * 1664525u + 1013904223u; q[i] = (uint8_t)(st >> 24); }
            k4_reencode(q, e); k4_restore(e, r);
            if (memcmp(q, r, 12)) return fail(B200_ERR_INVALID, "layout selftest: k4_reencode/k4_restore are not inverses");
        }
    }
    const int types[] = {T_Q8_0, T_Q4_K, T_Q5_K, T_Q6_K};
    for (int t : types) {
        const int gb = group_bytes(t), be = block_elems(t), bb = block_bytes(t), blocks_per_group = GROUP_ELEMS / be;
        for (int K : {256, 512, 2048, 4096, 14336 / 256 * 256}) {
            const int ng = n_groups(K); const int64_t rrb = (int64_t)raw_row_bytes(t, K);
            std::vector<int> hits((size_t)rrb, 0);
            for (int g = 0; g < ng; g++)
                for (int o = 0; o < gb; o++) {
                    int64_t s = src_byte(t, K, g, o);
                    if (s < 0) continue;
                    if (s >= rrb) return fail(B200_ERR_INVALID, "layout selftest: src_byte out of range");
                    hits[(size_t)s]++;
                    int blk_global = (int)(s / bb), i = (int)(s % bb);
                    if (blk_global / blocks_per_group != g) return fail(B200_ERR_INVALID, "layout selftest: group mismatch");
                    if (dst_byte(t, blk_global % blocks_per_group, i) != o) return fail(B200_ERR_INVALID, "layout selftest: dst_byte is not the inverse of src_byte");
                }
            for (int h : hits) if (h != 1) return fail(B200_ERR_INVALID, "layout selftest: raw byte not covered exactly once");
        }
        // tile layout of the dequant-in-GEMM path (qgemm.cuh): every blob byte names one raw byte of one of the 128 rows'
        // blocks of the K-step, and every such raw byte is named exactly once
        const int blob = qg_blob_bytes(t), blocks_per_step = QG_KSTEP / be;
        if (blob != QG_ROWS * blocks_per_step * bb) return fail(B200_ERR_INVALID, "layout selftest: tile blob size");
        std::vector<int> seen((size_t)blob, 0);
        for (int o = 0; o < blob; o++) {
            int row, blk, i; tile_src(t, o, row, blk, i);
            if (row < 0 || row >= QG_ROWS || blk < 0 || blk >= blocks_per_step || i < 0 || i >= bb) return fail(B200_ERR_INVALID, "layout selftest: tile_src out of range");
            seen[((size_t)row * blocks_per_step + blk) * bb + i]++;
        }
        for (int h : seen) if (h != 1) return fail(B200_ERR_INVALID, "layout selftest: tile layout does not cover every raw byte exactly once");
    }
    // stream-K arithmetic: the CTA ranges tile [0, T) in order, and qg_cta_of_step inverts qg_cta_first_step
    for (int G : {1, 7, 148}) for (int T : {1, 5, 147, 148, 149, 512, 768, 1792, 3584, 16032, 60000}) {
        if (qg_cta_first_step(0, T, G) != 0 || qg_cta_first_step(G, T, G) != T) return fail(B200_ERR_INVALID, "layout selftest: stream-K ranges do not cover [0, T)");
        for (int c = 0; c < G; c++) {
            const int b = qg_cta_first_step(c, T, G), e = qg_cta_first_step(c + 1, T, G);
            if (e < b) return fail(B200_ERR_INVALID, "layout selftest: stream-K range is not monotone");
            for (int st = b; st < e; st += std::max(1, (e - b) / 3)) if (qg_cta_of_step(st, T, G) != c) return fail(B200_ERR_INVALID, "layout selftest: qg_cta_of_step");
            if (e > b && qg_cta_of_step(e - 1, T, G) != c) return fail(B200_ERR_INVALID, "layout selftest: qg_cta_of_step at a range end");
        }
    }
    for (int k = 0; k < 64; k++) if (kperm4(kperm4(k)) != k || (kperm4(k) & ~3) != (k & ~3)) return fail(B200_ERR_INVALID, "layout selftest: kperm4");
    return B200_OK;
}

}  // namespace b200
