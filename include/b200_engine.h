/*
 * b200_engine.h — C ABI of the B200-native GGUF inference engine (libb200engine.so).
 *
 * This is the drop-in boundary for the reference's generation path.  The reference worker
 * (/root/reference/nats_llm_studio.go) reaches its backend through LMStudioClient (HTTP + `lms`
 * CLI); a maintainer replaces that client with a cgo binding of the functions below (see
 * INTEGRATION.md).  Conventions: plain pointers and sizes only; return 0 = ok, <0 = error
 * (b200_last_error() gives the thread-local message); no C++ exceptions cross the boundary; the
 * caller owns every input buffer; the engine owns all device memory; entry points are
 * thread-safe; nothing calls back into the host language (poll-based), so cgo never re-enters.
 */
#ifndef B200_ENGINE_H
#define B200_ENGINE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200_model b200_model;

enum { B200_OK = 0, B200_ERR_INVALID = -1, B200_ERR_CUDA = -2, B200_ERR_NOMEM = -3, B200_ERR_UNSUPPORTED = -4,
       B200_ERR_NOT_FOUND = -5, B200_ERR_MORE = -6, B200_ERR_STATE = -7 };

typedef struct b200_config {
    int device;          /* CUDA device ordinal */
    int max_ctx;         /* KV positions reserved per sequence (0 = 4096) */
    int max_seqs;        /* concurrent sequences the KV pool is sized for (0 = 1) */
    uint32_t kv_pages;   /* KV pages (16 positions each) in a model's pool; 0 = max_seqs x max_ctx / 16: every sequence slot fully backed.
                          * A smaller pool oversubscribes the slots: pages are handed out as sequences grow, returned when a request
                          * ends, and the scheduler admits a request only when the pages of its prompt + max_tokens are free. */
} b200_config;

/* ggml type ids used in the kernel-level entry points (gguf/constants.py GGMLQuantizationType) */
enum { B200_F32 = 0, B200_F16 = 1, B200_Q8_0 = 8, B200_Q4_K = 12, B200_Q5_K = 13, B200_Q6_K = 14 };

typedef struct b200_sampling {
    float temperature;   /* <= 0: greedy */
    int top_k;           /* 0: off */
    float top_p;         /* >= 1: off */
    uint64_t seed;
    int max_tokens;
    int stop_token;      /* -1: none (EOS from the GGUF is always honoured unless ignore_eos) */
    int ignore_eos;
    int timeout_ms;      /* > 0: the request is ended with B200_FINISH_TIMEOUT once it has been in the engine this long
                            (the reference bounds every handler with a context deadline, nats_llm_studio.go:328) */
} b200_sampling;

enum { B200_EV_TOKEN = 1, B200_EV_FINISH = 2, B200_EV_ERROR = 3 };
enum { B200_FINISH_STOP = 1, B200_FINISH_LENGTH = 2, B200_FINISH_CANCELLED = 3, B200_FINISH_TIMEOUT = 4 };
typedef struct b200_event {
    uint64_t req_id; int kind; int token; int finish_reason;
    int n_prompt; int n_generated; float ttft_ms; float decode_ms;
} b200_event;

/* ---- lifecycle.  Replaces NewLMStudioClient (nats_llm_studio.go:34-44): the backend is in-process. */
int b200_init(const b200_config* cfg);
int b200_shutdown(void);
const char* b200_last_error(void);
/* Pinned host staging for uploads (JetStream Object Store chunks land here; README.md:250-318). */
void* b200_pinned_alloc(size_t n);
void b200_pinned_free(void* p);

/* ---- model registry.  Replaces PullModel (`lms get`, nats_llm_studio.go:46-59), ListModels
 * (GET /api/v0/models, :136-156) and DeleteModel (`lms unload` + RemoveAll, :87-133): models are
 * HBM-resident; pull streams .gguf bytes host -> HBM, delete frees HBM. */
int b200_model_open(const char* id, const void* gguf_head, size_t n, b200_model** out);  /* B200_ERR_MORE: pass a longer prefix */
int b200_model_upload(b200_model* m, uint64_t file_off, const void* bytes, size_t n);
int b200_model_commit(b200_model* m);                    /* repack on GPU, build the decode schedule */
int b200_model_close(b200_model* m);                     /* frees HBM, removes from registry */
int b200_model_load_file(const char* id, const char* path, b200_model** out);   /* open+upload+commit */
int b200_model_find(const char* id, b200_model** out);
int b200_model_list(char* json_out, size_t cap, size_t* need);   /* {"data":[{id,publisher,...}],"object":"list"} */
int b200_model_info(const b200_model* m, char* json_out, size_t cap, size_t* need);

/* ---- generation.  Replaces LMStudioClient.Chat (POST /api/v0/chat/completions, :158-179). */
int b200_submit(b200_model* m, const int32_t* toks, int n, const b200_sampling* s, uint64_t req_id);
int b200_poll(b200_event* out, int max, int timeout_ms);          /* returns number of events */
int b200_cancel(uint64_t req_id);
/* Synchronous single-sequence forms (what submit/poll drive underneath; used by tests and bench). */
int b200_model_generate(b200_model* m, const int32_t* prompt, int n_prompt, int n_gen, const b200_sampling* s,
                        int32_t* out_tokens, int* n_out, float* ttft_ms, float* decode_ms);
int b200_model_eval(b200_model* m, const int32_t* toks, int n, float* logits_all /* [n][n_vocab] or NULL */,
                    float* logits_last /* [n_vocab] or NULL */);
int b200_model_n_vocab(const b200_model* m);
/* tokenizer of the loaded GGUF (byte-level for the synthetic vocab) */
int b200_tokenize(const b200_model* m, const char* text, size_t n, int32_t* out, int cap, int* n_out);
int b200_detokenize(const b200_model* m, const int32_t* toks, int n, char* out, size_t cap, size_t* need);

/* Host-only forms (no GPU): the tokenizer and the chat template of a GGUF header (the first bytes of a .gguf file, up to
 * the end of its metadata).  The reference forwards `messages` to a runtime that applies both (nats_llm_studio.go:158-167);
 * here: byte-level BPE from tokenizer.ggml.tokens/merges (Llama-3 / GPT-2 pre-tokeniser split) and the Llama-3, granite and
 * ChatML chat layouts, recognised from tokenizer.chat_template.  b200_gguf_render_chat takes the JSON array of
 * {role, content} messages and returns the prompt's token ids, ending with the assistant turn's header. */
int b200_gguf_tokenize(const void* gguf_head, size_t n, const char* text, size_t text_n, int parse_special, int32_t* out, int cap, int* n_out);
int b200_gguf_detokenize(const void* gguf_head, size_t n, const int32_t* toks, int n_toks, char* out, size_t cap, size_t* need);
int b200_gguf_render_chat(const void* gguf_head, size_t n, const char* messages_json, size_t json_n, int32_t* out, int cap, int* n_out);

/* ---- subject handlers: same payloads in, same {ok,error,data} envelope bytes out as
 * Server.onListModels / OnPullModel / OnDeleteModel / OnChatModel (nats_llm_studio.go:228-364).
 * subject is "lmstudio.list_models" | "lmstudio.pull_model" | "lmstudio.delete_model" |
 * "lmstudio.chat_model".  The reply is what msg.Respond would send. */
int b200_handle(const char* subject, const void* payload, size_t n, char* reply, size_t cap, size_t* need);
/* ---- the worker.  The reference describes the service (README.md:475-502: connect to NATS_URL, queue-subscribe to the
 * four subjects with NATS_QUEUE_GROUP, reply to every request) but holds no code for it (no main, no Subscribe call).
 * b200_worker_run is that service: a NATS core-protocol client that queue-subscribes to lmstudio.list_models / pull_model /
 * delete_model / chat_model (and lmstudio.sync_model_from_bucket, README.md:286) and answers each message on its reply
 * subject with the handlers' envelope.  NULL arguments take NATS_URL (default nats://127.0.0.1:4222) and NATS_QUEUE_GROUP
 * (default lmstudio-workers) from the environment; LMSTUDIO_MODELS_DIR is honoured.  Requests are handled on n_threads
 * threads (0 = 2 x max_seqs), so concurrent chat requests are batched by the engine.  Blocks until b200_worker_stop()
 * (returns 0) or until the connection drops (< 0). */
int b200_worker_run(const char* nats_url, const char* queue_group, int n_threads);
int b200_worker_stop(void);
/* where pull_model looks for "<identifier>.gguf" / "<identifier>" files (LMSTUDIO_MODELS_DIR) */
int b200_set_models_dir(const char* dir);

/* ---- kernel-level entry points (parity tests, roofline bench).  Host buffers in and out. */
/* repack raw ggml blocks on the GPU and dequantise them back: out[n_rows*K] fp32, bit-exact */
int b200_dequantize(int type, const void* raw_blocks, int64_t n_rows, int K, float* out);
/* No counterpart in the reference.  b200_dequantize for callers that must not allocate device memory: a worker
 * that has already handed every spare byte of HBM to resident models and KV pages (a cudaMalloc there can fail, or
 * fragment the pool), or code on the request path, where cudaMalloc/cudaFree serialise against everything running on
 * the device and would stall generations in flight.  d_raw (n_rows rows of raw ggml blocks) and d_out
 * (n_rows*K floats) are DEVICE pointers owned by the caller; the engine-layout intermediate lives in a fixed
 * B200_NOALLOC_SCRATCH_BYTES array built into the library (resident on every GPU that loads it until exit), so
 * type/n_rows/K whose intermediate is larger are refused with B200_ERR_INVALID.  Neither this call nor anything it
 * invokes calls cudaMalloc, cudaFree or another allocating routine; b200_init must have run.  The first call on a
 * device also looks up the array's device address (make that call before starting any stream capture).  Every call
 * reuses the same array, which keeps the previous call's bytes (harmless: each call overwrites what it reads);
 * calls from this library are serialised, but work that overlaps on the GPU from elsewhere on the same array races.
 * The work is queued on a utility stream that b200_init created (not the generation stream, which the scheduler may
 * be capturing into a CUDA graph) and the call returns after it completed: inputs written on another stream must be
 * complete before calling. */
#define B200_NOALLOC_SCRATCH_BYTES (4u << 20)
int b200_dequantize_noalloc(int type, const void* d_raw, int64_t n_rows, int K, float* d_out);
/* one draw of the sampler (top-k -> top-p at T=1 -> temperature -> categorical with Philox4x32-10(seed, step)) over host
 * logits[V]: what chat_model's `temperature` / `top_p` / `top_k` / `seed` fields drive (README.md:196-205) */
int b200_sample(const float* logits, int V, const b200_sampling* s, uint64_t step, int* token);
/* y[N] = W[N x K] . x[K] through the streaming GEMV; iters>1 repeats the launch, ms = mean per iter */
int b200_gemv(int type, const void* raw_blocks, int64_t N, int K, const float* x, float* y, int iters, float* ms);
/* C[M x N] = A[M x K] . B[N x K]^T through the tcgen05/TMEM/TMA GEMM that prompt ingestion uses (inputs rounded to
 * bf16, fp32 accumulate); K % 64 == 0.  Host buffers. */
int b200_gemm_bf16(const float* A, const float* B, float* C, int M, int N, int K);
/* y[B x N] = x[B x K] . W[N x K]^T through the dequant-in-GEMM kernel batched decode uses (weights stay quantised in HBM,
 * expanded to fp16 in shared memory, tcgen05 fp16 x fp16 -> fp32); raw ggml rows, B <= 64, K % 256 == 0.  Host buffers;
 * iters > 1 repeats the launch, ms = mean per iteration. */
int b200_qgemm(int type, const void* raw_blocks, int64_t N, int K, const float* x, int B, float* y, int iters, float* ms);
/* prompts with at least n tokens are ingested by the batched tensor-core path, shorter ones token by token through
 * the decode kernel (n <= 0: never batch).  Default 32. */
int b200_model_set_prefill_min(b200_model* m, int n);
/* Batched decode building blocks (what the b200_submit scheduler drives when max_seqs > 1): ingest a prompt prefix
 * into sequence slot `slot`'s KV pages, and advance n sequences by one greedy token each (row i: token toks[i] of
 * slot slots[i] at position pos[i]; logits, when not NULL, receives [n][n_vocab] fp32).  A row's result does not depend
 * on which other rows are in the batch. */
int b200_model_prefill_slot(b200_model* m, int slot, const int32_t* toks, int n, int pos0);
/* KV pages are allocated to a slot as its positions are written (B200_ERR_NOMEM when the pool is exhausted); a caller that
 * drives the slots itself returns a finished sequence's pages with b200_model_release_slot.  b200_model_kv_free_pages: pages
 * currently unassigned (< 0: error). */
int b200_model_release_slot(b200_model* m, int slot);
int b200_model_kv_free_pages(b200_model* m);
int b200_model_batch_step(b200_model* m, const int32_t* toks, const int32_t* slots, const int32_t* pos, int n, int32_t* next, float* logits);
/* The same step with per-row sampling (the `temperature`, `top_k`, `top_p`, `seed` request fields, README.md:196-205): row i
 * with samp[i].temperature > 0 draws its token with Philox(seed, draw[i]) where draw[i] = tokens that request has generated
 * so far — the counter the single-sequence path uses, so a seeded request yields the same tokens alone or in any batch. */
int b200_model_batch_step_sampled(b200_model* m, const int32_t* toks, const int32_t* slots, const int32_t* pos, int n,
                                  const b200_sampling* samp, const int32_t* draw, int32_t* next, float* logits);
/* ---- tensor parallelism.  No counterpart in the reference, whose only scale-out is replicas in a NATS
 * queue group (README.md:484).  One process per GPU; rank 0 creates a 128-byte NCCL id, the host side distributes it
 * (the Go worker would publish it on a NATS subject; tests and tools use torch.distributed), every rank then calls
 * b200_tp_init after b200_init and before opening a model.  A model opened afterwards keeps only this rank's shard
 * (attention heads, FFN columns, vocabulary rows: b200_tp_shard_window).  Every rank must then make the same calls with
 * the same arguments:
 *   - b200_model_generate / b200_model_eval / b200_bench_decode (greedy): the persistent batch-1 kernel runs on every
 *     rank; the partial sums of the two row-parallel products of a layer and the argmax candidates are exchanged INSIDE
 *     the kernel as tagged words written straight into the peers' memory over NVLink (buffers shared through CUDA IPC);
 *     b200_model_eval fills this rank's vocabulary columns of the logits array;
 *   - b200_model_prefill_slot / b200_model_batch_step (batched rows): partial sums all-reduced with NCCL, logits slices
 *     all-gathered. */
int b200_tp_unique_id(void* out, size_t cap);
int b200_tp_init(int rank, int world, const void* id, size_t n);
/* Host arithmetic only (no GPU): the part of a GGUF tensor [N rows x K] that rank `rank` of `world` keeps.  kind is
 * "attn_q" | "attn_k" | "attn_v" | "ffn_gate" | "ffn_up" | "output" (a row range, all of K) or "attn_output" | "ffn_down"
 * (all rows, a K window aligned to 256-element blocks). */
int b200_tp_shard_window(const char* kind, int64_t N, int K, int rank, int world, int64_t* row0, int64_t* nrows, int* col0, int* ncols);
/* decode-step timing: n_steps greedy tokens from the current position of a primed context */
int b200_bench_decode(b200_model* m, int ctx_len, int n_steps, int warmup, float* ms_per_token, double* bytes_per_token);
/* batched-step timing (BASELINE.json metric "bs64"; config 4's bs1/bs32 on a tensor-parallel engine): n_seqs sequences at
 * context ctx advance by one token per step, `iters` steps between two CUDA events on the engine stream (no host copies
 * inside); bytes_per_step = algorithmic HBM bytes of one step on this GPU (weights once, n_seqs embedding rows, K/V of
 * every sequence) */
int b200_bench_batch_step(b200_model* m, int n_seqs, int ctx, int iters, int warmup, float* ms_per_step, double* bytes_per_step);
/* number of kernels this library has launched so far (bench.py reports the delta as gpu_launches) */
uint64_t b200_launch_count(void);
/* layout self-test (host only, no GPU): src_byte/dst_byte are inverse bijections for every type */
int b200_layout_selftest(void);

#ifdef __cplusplus
}
#endif
#endif
