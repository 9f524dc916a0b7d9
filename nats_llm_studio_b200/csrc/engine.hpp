// engine.hpp — host-side declarations shared by engine.cu, handlers.cpp and capi.cpp.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <chrono>
#include <condition_variable>
#include <cstdint>
#include <deque>
#include <map>
#include <memory>
#include <atomic>
#include <mutex>
#include <set>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../../include/b200_engine.h"
#include "gguf.hpp"
#include "tokenizer.hpp"

namespace b200 {

struct DecodeParams;
struct Phase;

extern thread_local std::string g_last_error;
int fail(int code, const std::string& msg);

struct Matrix { uint8_t* w = nullptr; int type = 0; int64_t N = 0, n_rows = 0; int K = 0;
                uint8_t* tl = nullptr; };   // tl: tile-layout copy for the dequant-in-GEMM path (qgemm.cuh), built on first batched use
struct Layer { float* attn_norm = nullptr; float* ffn_norm = nullptr; Matrix wq, wk, wv, wo, wgu, wd;
               __nv_bfloat16 *b_qkv = nullptr, *b_o = nullptr, *b_gu = nullptr, *b_d = nullptr; };   // bf16 copies (batched decode)

struct Engine {
    std::mutex mu;
    bool ready = false;
    int device = 0, max_ctx = 4096, max_seqs = 1, n_sm = 0, smem_per_sm = 0, smem_optin = 0;
    int kv_pages = 0;                            // b200_config.kv_pages: pages of 16 positions in a model's KV pool (0 = every slot fully backed)
    cudaStream_t stream = nullptr, copy_stream = nullptr;
    cudaStream_t aux_stream = nullptr;           // allocation-free utility calls (never captured into a step graph)
    uint64_t launches = 0;                       // kernels launched by this library (bench.py reports it)
    float* qg_ws = nullptr;                      // stream-K partial tiles of the dequant-in-GEMM kernel
    // tensor parallelism: one process per GPU, NCCL communicator created from an id the host side distributes
    int tp_rank = 0, tp_world = 1; void* nccl_comm = nullptr;
    int tp_unique_id(void* out, size_t cap);
    int tp_init(int rank, int world, const void* id, size_t n);
    int all_reduce_sum(float* buf, size_t n);                            // in place, on `stream`
    int all_gather(const float* src, float* dst, size_t n_per_rank);     // dst = [world][n_per_rank], on `stream`
    int init(const b200_config* cfg);
    int ensure();
    void shutdown();                             // after every model is closed: a later init() may use a different config
};
Engine& engine();

struct Model {
    std::string id, arch;
    GgufMeta meta;
    int n_layer = 0, n_embd = 0, n_head = 0, n_head_kv = 0, head_dim = 0, n_ff = 0, n_vocab = 0, n_rot = 0, n_ctx_train = 0;
    float rms_eps = 1e-5f, rope_base = 10000.f, emb_scale = 0, res_scale = 0, attn_scale = 0, logit_scale = 0;
    int bos = -1, eos = -1;
    std::vector<std::string> vocab; std::vector<int> tok_type;
    // tokenizer tables
    Tokenizer tok;
    // upload state
    uint64_t uploaded = 0; bool committed = false;
    // staging ring (uploads land here tensor by tensor; see Model::open)
    struct StageSlot { uint64_t begin, end; cudaEvent_t freed; bool recorded; int tensor; };
    uint8_t* d_stage = nullptr; uint64_t stage_bytes = 0, stage_head = 0, next_off = 0; std::deque<StageSlot> stage_live;
    std::vector<int> t_order; std::vector<uint64_t> t_got, t_ring; size_t t_cursor = 0; std::set<std::string> t_done; std::vector<float> rope_ff;
    int g_head = 0, g_head_kv = 0, g_ff = 0;     // whole-model head / FFN counts (n_head, n_head_kv, n_ff are this rank's)
    int stage_alloc(uint64_t nbytes, uint64_t* off);
    int upload_async(uint64_t file_off, const void* bytes, size_t n, cudaEvent_t* ticket);
    int process_tensor(const GgufTensor& t, const uint8_t* src);
    void release_staging();
    uint64_t weight_bytes = 0, hbm_bytes = 0;
    std::vector<void*> owned;
    // weights
    Matrix tok_embd, out_w; float* out_norm = nullptr; float* inv_freq = nullptr; std::vector<Layer> layers;
    // runtime
    Phase* d_phases = nullptr; int n_phases = 0;
    float* d_logits = nullptr; uint32_t* d_ll16 = nullptr; uint2* d_ll32 = nullptr; int hown_cap = 0;
    int *d_state = nullptr, *d_out_tokens = nullptr, *d_forced = nullptr, *d_block_table = nullptr;
    __half* d_kv = nullptr; int n_pages = 0, n_splits = 1;
    // ---- KV page allocator (K5): the pool's pages are handed to sequence slots on demand and returned when a request ends;
    // the device block table [slot][page index] is what every attention kernel reads (no kernel assumes an identity map).
    std::mutex kv_mu; std::vector<int> kv_free; std::vector<std::vector<int>> kv_slot_pages; std::vector<int> kv_host_table;
    int kv_ensure(int slot, int upto_pos);       // back positions [0, upto_pos] of `slot`; new table entries are uploaded on the engine stream
    void kv_release(int slot);                   // return the slot's pages (stream order keeps kernels already enqueued safe)
    int kv_free_count();
    int sync_slot = 0;                           // KV sequence slot of the synchronous batch-1 path (generate / eval / chat with max_seqs == 1);
                                                 // with max_seqs > 1 it is an extra slot the scheduler never hands out
    int max_segs = 0;
    int grid = 0, smem_max_grp = 0; uint32_t ring_bytes = 0; size_t smem_bytes = 0; const void* decode_fn = nullptr;   // decode kernel variant of this model
    int cur_pos = 0;
    long long* d_trace = nullptr;
    // prefill scratch (allocated on first use)
    static constexpr int PREFILL_CHUNK = 2048;
    float *pf_H = nullptr, *pf_Y = nullptr; __nv_bfloat16 *pf_X = nullptr, *pf_Q = nullptr, *pf_ACT = nullptr, *pf_W = nullptr; int* pf_toks = nullptr;
    float* bt_logits = nullptr; int* bt_next = nullptr; struct RowSampling* bt_samp = nullptr; __nv_bfloat16* b_out = nullptr; bool bf16_ready = false; int pages_per_seq = 0;
    int prefill_min = 32;                        // prompts shorter than this go through the decode kernel token by token
    std::mutex run_mu;
    std::atomic<int> lock_waiters{0};            // threads that want run_mu to release the model: the scheduler steps aside for them (std::mutex is not fair)
    void release_locked() { lock_waiters++; { std::lock_guard<std::mutex> lk(run_mu); release(); } lock_waiters--; }

    ~Model();
    void release();
    template <typename T> int own(T** p, size_t n);
    int open(const std::string& id, const uint8_t* head, size_t n);
    int upload(uint64_t file_off, const void* bytes, size_t n);
    int commit();
    void fill_params(DecodeParams& P) const;
    int run_steps(int n_steps, const int32_t* forced, int n_forced, bool reset, float* logits_all_dev);
    int generate(const int32_t* prompt, int n_prompt, int n_gen, const b200_sampling* s, int32_t* out, int* n_out, float* ttft_ms, float* decode_ms,
                 int* finish_reason = nullptr, const std::chrono::steady_clock::time_point* t_submit = nullptr);   // t_submit: start of s->timeout_ms
    int eval(const int32_t* toks, int n, float* logits_all, float* logits_last);
    double bytes_per_token(int ctx) const;
    int bench_decode(int ctx_len, int n_steps, int warmup, float* ms_per_token, double* bytes);
    int ensure_prefill_buffers();
    int ensure_bf16_weights();
    int ensure_tile_weights();
    int tp_reduce_add(int M, const struct QgSplit& sp);
    bool tp = false; int n_vocab_l = 0; float* pf_T = nullptr;    // tensor-parallel shard: local vocab rows, partial-sum buffer
    // tensor parallelism inside the batch-1 kernel: exchange buffers in peer memory (this rank's allocation + the peers' IPC mappings)
    uint8_t* tp_buf = nullptr; void* tp_peer[8] = {}; size_t tp_rx_bytes = 0; unsigned* tp_error = nullptr;
    int tp_setup_exchange();
    int tp_check_error();
    bool use_qgemm(int M) const;
    bool tile_ready = false;
    struct RowRun { int row0, n, pos0, slot; };      // rows [row0, row0+n) of a chunk = positions pos0.. of sequence `slot`
    int dense_rows(const int* d_toks, int M, const int* d_pos, const int* d_slot, int pos0, int slot0, bool want_logits, bool q, const std::vector<RowRun>* runs = nullptr);
    int ensure_attn_maps(); void* tmap_q = nullptr; void* tmap_kv = nullptr; void* tmap_qd = nullptr; std::vector<uint8_t> tmap_storage;
    int prefill_rows(const int32_t* toks, const int32_t* slots, const int32_t* pos, int M);
    void maybe_bf16_cache();
    bool bf16_tried = false;
    std::vector<cudaGraphExec_t> step_graph; std::vector<char> step_seen; std::vector<int> step_launches; bool graph_failed = false;   // batch_step: one CUDA graph per batch size
    int prefill_slot(int slot, const int32_t* toks, int n, int pos0);
    int prefill(const int32_t* toks, int n);
    int batch_step(const int32_t* toks, const int32_t* slots, const int32_t* pos, int B, int32_t* next, float* logits = nullptr,
                   const b200_sampling* samp = nullptr, const int32_t* draw = nullptr);   // samp[B], draw[B] (tokens already generated per row) or null = greedy
    int step_enqueue(int B);
    int bench_batch(int B, int ctx, int iters, int warmup, float* ms_per_step, double* bytes);
    // tokenizer (tokenizer.cpp)
    void build_tokenizer();
    std::vector<int32_t> tokenize(const std::string& text, bool parse_special) const;
    std::string detokenize(const int32_t* toks, int n) const;
    std::string info_json() const;
};

int sample_token(Engine& E, Model& m, const b200_sampling& s, uint64_t step);
int kernel_sample(const float* logits, int V, const b200_sampling& s, uint64_t step, int* token);
int kernel_dequantize(int type, const void* raw, int64_t n_rows, int K, float* out);
int kernel_gemv(int type, const void* raw, int64_t N, int K, const float* x, float* y, int iters, float* ms);
int kernel_gemm_bf16(const float* A, const float* B, float* C, int M, int N, int K);
int kernel_qgemm(int type, const void* raw, int64_t N, int K, const float* x, int B, float* y, int iters, float* ms);
int kernel_dequantize_noalloc(int type, const void* d_raw, int64_t n_rows, int K, float* d_out);
int layout_selftest();
// rows / K window of a tensor that tensor-parallel rank `rank` of `world` keeps (host arithmetic only; Model::commit uses it)
int tp_shard_window(const char* kind, int64_t N, int K, int rank, int world, int64_t* row0, int64_t* nrows, int* col0, int* ncols);

// registry (handlers.cpp)
struct Registry {
    std::mutex mu;
    std::map<std::string, std::shared_ptr<Model>> models;
    std::string models_dir;
};
Registry& registry();
std::string handle_subject(const std::string& subject, const std::string& payload);
// tokenizer of a GGUF header (host only): tokenizer.ggml.* -> Tokenizer; "" or an error message
std::string tokenizer_from_gguf(const GgufMeta& meta, int n_vocab, Tokenizer& out);
// chat rendering (handlers.cpp): messages (parsed JSON array as text) -> prompt token ids with the model's template
std::string render_chat(const Tokenizer& tok, const std::string& messages_json, std::vector<int32_t>& toks);
// NATS worker (worker.cpp): queue-subscribes to the subjects and answers through handle_subject until stopped
int worker_run(const char* url, const char* queue_group, int n_threads);
int worker_stop();
// blocking generation through the continuous-batching scheduler (capi.cpp): what lmstudio.chat_model uses when max_seqs > 1
int sched_generate(const std::shared_ptr<Model>& m, const std::vector<int32_t>& prompt, const b200_sampling& s, std::vector<int32_t>& out, float* ttft_ms, float* decode_ms, int* finish_reason);

}  // namespace b200
