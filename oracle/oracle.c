/*
 * oracle.c — CPU restatement of the chat_model generation path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the parity checker for the B200 engine.  It is never linked into, imported
 * by, or called from the product path (nats_llm_studio_b200/); only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * What it restates
 * ----------------
 * The reference (`/root/reference/nats_llm_studio.go:158-179`, `LMStudioClient.Chat`) performs
 * no arithmetic itself: it POSTs the request to LM Studio, whose bundled llama.cpp runtime
 * (closed, un-vendored, version un-pinned: scripts/setup_unix.sh:14-32) does the work.  The
 * algorithm therefore lives in a third-party dependency that is absent from /root/reference.
 * This file restates its published algorithm:
 *   - GGUF v3 container + ggml block formats, cross-checked bit-for-bit on this image against
 *     gguf-py 0.19.0 (gguf/quants.py:378-401 Q8_0, :475-522 Q4_K, :525-549 Q5_K, :552-572 Q6_K;
 *     gguf/constants.py:10-12 magic/version/alignment, :4241-4246 block sizes);
 *   - the standard Llama / Granite decoder forward pass in fp32 with fp64 accumulation
 *     (RMSNorm -> QKV -> RoPE(adjacent pairs, "NORM" type) -> GQA attention -> wo -> residual
 *      -> RMSNorm -> SwiGLU -> residual; final norm -> lm_head), with Granite's four scalars
 *     (gguf/constants.py:132,142,143,176).
 *
 * PARITY PINNING: the dequantisers are pinned bit-exactly against gguf-py (tests/test_oracle.py).
 * The reference holds no golden vectors for the forward pass (it has zero tests, SURVEY.md
 * §4/§8c) and its runtime cannot be installed offline, so the forward pass is pinned to outputs
 * of an imported implementation instead: Hugging Face transformers' LlamaForCausalLM /
 * GraniteForCausalLM (CPU, float64) run on the gguf-py-dequantised tensors of the committed
 * GGUFs (tests/golden/make_hf_golden.py -> tests/golden/*.hf.npz; measured distance 2e-7 of the
 * largest logit), and additionally to an independent numpy forward (tests/golden/make_golden.py).
 *
 * Build: see oracle/Makefile (gcc -O2 -fopenmp -ffp-contract=off -shared -fPIC).
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <fcntl.h>
#include <unistd.h>
#include <sys/mman.h>
#include <sys/stat.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---------------------------------------------------------------- ggml types */
enum { T_F32 = 0, T_F16 = 1, T_Q8_0 = 8, T_Q4_K = 12, T_Q5_K = 13, T_Q6_K = 14, T_BF16 = 30 };
#define QK_K 256

static int type_block_elems(int t) {
    switch (t) { case T_F32: case T_F16: case T_BF16: return 1; case T_Q8_0: return 32;
                 case T_Q4_K: case T_Q5_K: case T_Q6_K: return QK_K; default: return 0; }
}
static int type_block_bytes(int t) {
    switch (t) { case T_F32: return 4; case T_F16: case T_BF16: return 2; case T_Q8_0: return 34;
                 case T_Q4_K: return 144; case T_Q5_K: return 176; case T_Q6_K: return 210; default: return 0; }
}

/* IEEE binary16 -> binary32, exact (subnormals, inf, nan payload kept). */
static float fp16_to_fp32(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu, man = h & 0x3ffu, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else { /* subnormal: normalise */
            int e = -1; do { e++; man <<= 1; } while (!(man & 0x400u));
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) bits = sign | 0x7f800000u | (man << 13);
    else bits = sign | ((exp + 112u) << 23) | (man << 13);
    float f; memcpy(&f, &bits, 4); return f;
}
static float bf16_to_fp32(uint16_t h) { uint32_t b = (uint32_t)h << 16; float f; memcpy(&f, &b, 4); return f; }
static uint16_t rd16(const uint8_t *p) { uint16_t v; memcpy(&v, p, 2); return v; }

/* K-quant 6-bit scale/min unpack — gguf/quants.py:480-503 (Q4_K.get_scale_min). */
static void get_scale_min_k4(int j, const uint8_t *q, uint8_t *sc, uint8_t *m) {
    if (j < 4) { *sc = q[j] & 63; *m = q[j + 4] & 63; }
    else { *sc = (q[j + 4] & 0x0F) | ((q[j - 4] >> 6) << 4); *m = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

/* Dequantise one block-row segment.  Float op order mirrors gguf-py exactly; the file is built
 * with -ffp-contract=off so `a*b - c` is two roundings as numpy performs it. */
static void dequant_q8_0(const uint8_t *b, int64_t nblk, float *y) {           /* quants.py:395-401 */
    for (int64_t i = 0; i < nblk; i++, b += 34, y += 32) {
        float d = fp16_to_fp32(rd16(b));
        for (int j = 0; j < 32; j++) y[j] = (float)(int8_t)b[2 + j] * d;
    }
}
static void dequant_q4_k(const uint8_t *b, int64_t nblk, float *y) {           /* quants.py:505-522 */
    for (int64_t i = 0; i < nblk; i++, b += 144, y += QK_K) {
        float d = fp16_to_fp32(rd16(b)), dmin = fp16_to_fp32(rd16(b + 2));
        const uint8_t *scales = b + 4, *qs = b + 16;
        for (int j = 0; j < 8; j++) {
            uint8_t sc, m; get_scale_min_k4(j, scales, &sc, &m);
            float dl = d * (float)sc, ml = dmin * (float)m;
            const uint8_t *q = qs + 32 * (j / 2); int sh = 4 * (j & 1);
            for (int l = 0; l < 32; l++) y[32 * j + l] = dl * (float)((q[l] >> sh) & 0xF) - ml;
        }
    }
}
static void dequant_q5_k(const uint8_t *b, int64_t nblk, float *y) {           /* quants.py:527-549 */
    for (int64_t i = 0; i < nblk; i++, b += 176, y += QK_K) {
        float d = fp16_to_fp32(rd16(b)), dmin = fp16_to_fp32(rd16(b + 2));
        const uint8_t *scales = b + 4, *qh = b + 16, *qs = b + 48;
        for (int j = 0; j < 8; j++) {
            uint8_t sc, m; get_scale_min_k4(j, scales, &sc, &m);
            float dl = d * (float)sc, ml = dmin * (float)m;
            const uint8_t *q = qs + 32 * (j / 2); int sh = 4 * (j & 1);
            for (int l = 0; l < 32; l++) {
                int v = ((q[l] >> sh) & 0xF) | (((qh[l] >> j) & 1) << 4);
                y[32 * j + l] = dl * (float)v - ml;
            }
        }
    }
}
static void dequant_q6_k(const uint8_t *b, int64_t nblk, float *y) {           /* quants.py:554-572 */
    for (int64_t i = 0; i < nblk; i++, b += 210, y += QK_K) {
        const uint8_t *ql = b, *qh = b + 128; const int8_t *sc = (const int8_t *)(b + 192);
        float d = fp16_to_fp32(rd16(b + 208));
        for (int e = 0; e < QK_K; e++) {
            int half = e / 128, r = e % 128, grp = r / 32, l = r % 32;
            uint8_t qlb = ql[64 * half + 32 * (grp & 1) + l];
            int lo = (grp < 2) ? (qlb & 0xF) : (qlb >> 4);
            int hi = (qh[32 * half + l] >> (2 * grp)) & 3;
            int q = (int)(int8_t)(lo | (hi << 4)) - 32;
            float dl = d * (float)sc[e / 16];
            y[e] = dl * (float)q;
        }
    }
}

/* Public: dequantise n_elems (multiple of the block size) of `type`.  Returns 0 or -1. */
int orc_dequantize(int type, const void *blocks, int64_t n_elems, float *out) {
    int be = type_block_elems(type);
    if (!be || n_elems % be) return -1;
    int64_t nb = n_elems / be; const uint8_t *b = (const uint8_t *)blocks;
    switch (type) {
        case T_F32: memcpy(out, b, (size_t)n_elems * 4); return 0;
        case T_F16: for (int64_t i = 0; i < n_elems; i++) out[i] = fp16_to_fp32(rd16(b + 2 * i)); return 0;
        case T_BF16: for (int64_t i = 0; i < n_elems; i++) out[i] = bf16_to_fp32(rd16(b + 2 * i)); return 0;
        case T_Q8_0: dequant_q8_0(b, nb, out); return 0;
        case T_Q4_K: dequant_q4_k(b, nb, out); return 0;
        case T_Q5_K: dequant_q5_k(b, nb, out); return 0;
        case T_Q6_K: dequant_q6_k(b, nb, out); return 0;
    }
    return -1;
}
int orc_type_block_elems(int t) { return type_block_elems(t); }
int orc_type_block_bytes(int t) { return type_block_bytes(t); }

/* ---------------------------------------------------------------- GGUF v3 reader */
enum { GV_U8, GV_I8, GV_U16, GV_I16, GV_U32, GV_I32, GV_F32, GV_BOOL, GV_STR, GV_ARR, GV_U64, GV_I64, GV_F64 };
static const int gv_size[] = { 1, 1, 2, 2, 4, 4, 4, 1, 0, 0, 8, 8, 8 };

typedef struct { const char *name; int name_len; int type; int n_dims; int64_t ne[4]; const uint8_t *data; } orc_tensor;
typedef struct { const char *key; int key_len; int type; const uint8_t *val; int arr_type; uint64_t arr_n; } orc_kv;

typedef struct orc_model {
    uint8_t *map; size_t map_len;
    orc_kv *kv; uint64_t n_kv;
    orc_tensor *tens; uint64_t n_tens;
    char arch[64];
    int n_layer, n_embd, n_head, n_head_kv, head_dim, n_ff, n_vocab, n_rot, n_ctx;
    float rms_eps, rope_base, emb_scale, res_scale, attn_scale, logit_scale;
    /* KV cache (fp32) */
    int kv_cap, n_past; float *kc, *vc;
    int act_quant;   /* 0: fp32 activations.  1: Q8 (per-256 block) activations, the engine's GEMV numerics */
    int kv_f16;      /* 1: round cached K/V through binary16 like the engine's KV cache */
    int xchg_f16;    /* 1: round the vectors the engine exchanges between SMs (hidden state as seen by the norms,
                        attention output, SwiGLU output) through binary16, as the engine's tagged fp16 words do */
} orc_model;

typedef struct { const uint8_t *p, *end; int err; } rdr;
static uint64_t r_u64(rdr *r) { if (r->p + 8 > r->end) { r->err = 1; return 0; } uint64_t v; memcpy(&v, r->p, 8); r->p += 8; return v; }
static uint32_t r_u32(rdr *r) { if (r->p + 4 > r->end) { r->err = 1; return 0; } uint32_t v; memcpy(&v, r->p, 4); r->p += 4; return v; }
static const char *r_str(rdr *r, int *len) {
    uint64_t n = r_u64(r); if (r->err || (uint64_t)(r->end - r->p) < n) { r->err = 1; *len = 0; return NULL; }
    const char *s = (const char *)r->p; r->p += n; *len = (int)n; return s;
}
static void r_skip_val(rdr *r, int type) {
    if (type == GV_STR) { int l; r_str(r, &l); }
    else if (type == GV_ARR) {
        int et = (int)r_u32(r); uint64_t n = r_u64(r);
        if (et == GV_STR) { for (uint64_t i = 0; i < n && !r->err; i++) { int l; r_str(r, &l); } }
        else if (et < 13 && gv_size[et]) { if ((uint64_t)(r->end - r->p) < n * gv_size[et]) r->err = 1; else r->p += n * gv_size[et]; }
        else r->err = 1;
    } else if (type < 13 && gv_size[type]) { if (r->p + gv_size[type] > r->end) r->err = 1; else r->p += gv_size[type]; }
    else r->err = 1;
}

static const orc_kv *find_kv(const orc_model *m, const char *key) {
    int n = (int)strlen(key);
    for (uint64_t i = 0; i < m->n_kv; i++) if (m->kv[i].key_len == n && !memcmp(m->kv[i].key, key, n)) return &m->kv[i];
    return NULL;
}
static double kv_num(const orc_model *m, const char *key, double dflt) {
    const orc_kv *k = find_kv(m, key); if (!k) return dflt;
    switch (k->type) {
        case GV_U8: case GV_BOOL: return *(uint8_t *)k->val; case GV_I8: return *(int8_t *)k->val;
        case GV_U16: { uint16_t v; memcpy(&v, k->val, 2); return v; } case GV_I16: { int16_t v; memcpy(&v, k->val, 2); return v; }
        case GV_U32: { uint32_t v; memcpy(&v, k->val, 4); return v; } case GV_I32: { int32_t v; memcpy(&v, k->val, 4); return v; }
        case GV_F32: { float v; memcpy(&v, k->val, 4); return v; }
        case GV_U64: { uint64_t v; memcpy(&v, k->val, 8); return (double)v; } case GV_I64: { int64_t v; memcpy(&v, k->val, 8); return (double)v; }
        case GV_F64: { double v; memcpy(&v, k->val, 8); return v; }
    }
    return dflt;
}
static double arch_num(const orc_model *m, const char *suffix, double dflt) {
    char key[160]; snprintf(key, sizeof key, "%s.%s", m->arch, suffix); return kv_num(m, key, dflt);
}
static const orc_tensor *find_tensor(const orc_model *m, const char *name) {
    int n = (int)strlen(name);
    for (uint64_t i = 0; i < m->n_tens; i++) if (m->tens[i].name_len == n && !memcmp(m->tens[i].name, name, n)) return &m->tens[i];
    return NULL;
}

void orc_free(orc_model *m) {
    if (!m) return;
    if (m->map) munmap(m->map, m->map_len);
    free(m->kv); free(m->tens); free(m->kc); free(m->vc); free(m);
}

orc_model *orc_load(const char *path) {
    int fd = open(path, O_RDONLY); if (fd < 0) return NULL;
    struct stat st; if (fstat(fd, &st)) { close(fd); return NULL; }
    orc_model *m = (orc_model *)calloc(1, sizeof *m);
    m->map_len = (size_t)st.st_size;
    m->map = (uint8_t *)mmap(NULL, m->map_len, PROT_READ, MAP_PRIVATE, fd, 0); close(fd);
    if (m->map == MAP_FAILED) { m->map = NULL; orc_free(m); return NULL; }
    rdr r = { m->map, m->map + m->map_len, 0 };
    if (r_u32(&r) != 0x46554747u) { orc_free(m); return NULL; }          /* "GGUF", constants.py:10 */
    uint32_t ver = r_u32(&r); if (ver != 3 && ver != 2) { orc_free(m); return NULL; }
    m->n_tens = r_u64(&r); m->n_kv = r_u64(&r);
    if (r.err || m->n_tens > (1u << 20) || m->n_kv > (1u << 20)) { orc_free(m); return NULL; }
    m->kv = (orc_kv *)calloc(m->n_kv ? m->n_kv : 1, sizeof(orc_kv));
    m->tens = (orc_tensor *)calloc(m->n_tens ? m->n_tens : 1, sizeof(orc_tensor));
    for (uint64_t i = 0; i < m->n_kv && !r.err; i++) {
        orc_kv *k = &m->kv[i]; k->key = r_str(&r, &k->key_len); k->type = (int)r_u32(&r); k->val = r.p;
        if (k->type == GV_ARR && r.p + 12 <= r.end) { uint32_t et; memcpy(&et, r.p, 4); k->arr_type = (int)et; memcpy(&k->arr_n, r.p + 4, 8); }
        r_skip_val(&r, k->type);
    }
    for (uint64_t i = 0; i < m->n_tens && !r.err; i++) {
        orc_tensor *t = &m->tens[i]; t->name = r_str(&r, &t->name_len); t->n_dims = (int)r_u32(&r);
        if (t->n_dims > 4) { r.err = 1; break; }
        for (int d = 0; d < 4; d++) t->ne[d] = 1;
        for (int d = 0; d < t->n_dims; d++) t->ne[d] = (int64_t)r_u64(&r);
        t->type = (int)r_u32(&r); t->data = (const uint8_t *)(uintptr_t)r_u64(&r); /* offset for now */
    }
    if (r.err) { orc_free(m); return NULL; }
    uint64_t align = (uint64_t)kv_num(m, "general.alignment", 32);           /* constants.py:12 */
    uint64_t base = (uint64_t)(r.p - m->map); base = (base + align - 1) / align * align;
    for (uint64_t i = 0; i < m->n_tens; i++) {
        orc_tensor *t = &m->tens[i]; uint64_t off = (uint64_t)(uintptr_t)t->data;
        int be = type_block_elems(t->type);
        if (!be) { orc_free(m); return NULL; }
        uint64_t nel = 1; for (int d = 0; d < t->n_dims; d++) nel *= (uint64_t)t->ne[d];
        uint64_t nbytes = nel / be * type_block_bytes(t->type);
        if (base + off + nbytes > m->map_len) { orc_free(m); return NULL; }
        t->data = m->map + base + off;
    }
    const orc_kv *ak = find_kv(m, "general.architecture");
    if (!ak || ak->type != GV_STR) { orc_free(m); return NULL; }
    { uint64_t n; memcpy(&n, ak->val, 8); if (n >= sizeof m->arch) n = sizeof m->arch - 1; memcpy(m->arch, ak->val + 8, n); m->arch[n] = 0; }
    m->n_layer = (int)arch_num(m, "block_count", 0);
    m->n_embd = (int)arch_num(m, "embedding_length", 0);
    m->n_head = (int)arch_num(m, "attention.head_count", 0);
    m->n_head_kv = (int)arch_num(m, "attention.head_count_kv", m->n_head);
    m->n_ff = (int)arch_num(m, "feed_forward_length", 0);
    m->n_ctx = (int)arch_num(m, "context_length", 4096);
    m->rms_eps = (float)arch_num(m, "attention.layer_norm_rms_epsilon", 1e-5);
    m->rope_base = (float)arch_num(m, "rope.freq_base", 10000.0);
    m->head_dim = m->n_head ? (int)arch_num(m, "attention.key_length", m->n_embd / m->n_head) : 0;
    m->n_rot = (int)arch_num(m, "rope.dimension_count", m->head_dim);
    m->emb_scale = (float)arch_num(m, "embedding_scale", 0.0);
    m->res_scale = (float)arch_num(m, "residual_scale", 0.0);
    m->attn_scale = (float)arch_num(m, "attention.scale", 0.0);
    m->logit_scale = (float)arch_num(m, "logit_scale", 0.0);
    const orc_tensor *te = find_tensor(m, "token_embd.weight");
    if (!te || !m->n_layer || !m->n_embd || !m->n_head) { orc_free(m); return NULL; }
    m->n_vocab = (int)te->ne[1];
    return m;
}

int orc_n_vocab(const orc_model *m) { return m->n_vocab; }
int orc_n_embd(const orc_model *m) { return m->n_embd; }
int orc_n_layer(const orc_model *m) { return m->n_layer; }
int orc_n_past(const orc_model *m) { return m->n_past; }
void orc_set_act_quant(orc_model *m, int mode) { m->act_quant = mode; }
void orc_set_kv_f16(orc_model *m, int on) { m->kv_f16 = on; }
void orc_set_xchg_f16(orc_model *m, int on) { m->xchg_f16 = on; }
void orc_set_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
int orc_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* Copy tensor (dequantised) into out; returns element count or -1. */
int64_t orc_tensor_f32(const orc_model *m, const char *name, float *out, int64_t cap) {
    const orc_tensor *t = find_tensor(m, name); if (!t) return -1;
    int64_t nel = 1; for (int d = 0; d < t->n_dims; d++) nel *= t->ne[d];
    if (!out) return nel;
    if (cap < nel) return -1;
    return orc_dequantize(t->type, t->data, nel, out) ? -1 : nel;
}

/* ---------------------------------------------------------------- forward pass */
static uint16_t fp32_to_fp16(float f) {   /* round-to-nearest-even */
    uint32_t x; memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u; int32_t e = (int32_t)((x >> 23) & 0xff) - 127 + 15; uint32_t man = x & 0x7fffffu;
    if (((x >> 23) & 0xff) == 0xff) return (uint16_t)(sign | 0x7c00u | (man ? 0x200u | (man >> 13) : 0));
    if (e >= 31) return (uint16_t)(sign | 0x7c00u);
    if (e <= 0) {
        if (e < -10) return (uint16_t)sign;
        man |= 0x800000u; int shift = 14 - e; uint32_t hm = man >> shift, rem = man & ((1u << shift) - 1), halfway = 1u << (shift - 1);
        if (rem > halfway || (rem == halfway && (hm & 1))) hm++;
        return (uint16_t)(sign | hm);
    }
    uint32_t hm = man >> 13, rem = man & 0x1fffu; uint16_t h = (uint16_t)(sign | ((uint32_t)e << 10) | hm);
    if (rem > 0x1000u || (rem == 0x1000u && (hm & 1))) h++;
    return h;
}

/* Engine-numerics mode (act_quant=1): activations are quantised per 256-element block to int8
 * with an fp32 scale d = amax/127 (q = round-half-away(x/d)), the dot product over each block is
 * taken on the dequantised weights in double.  This isolates activation rounding from ordering. */
static void quantise_act_256(const float *x, int n, float *xq) {
    for (int b = 0; b < n; b += 256) {
        int e = b + 256 < n ? b + 256 : n; float amax = 0;
        for (int i = b; i < e; i++) { float a = fabsf(x[i]); if (a > amax) amax = a; }
        float d = amax / 127.0f, id = amax > 0 ? 127.0f / amax : 0.0f;
        for (int i = b; i < e; i++) { float v = x[i] * id; float q = nearbyintf(v); if (q > 127.0f) q = 127.0f; if (q < -127.0f) q = -127.0f; xq[i] = q * d; }
    }
}

/* y[N] = W[N x K] . x[K], W any supported type, dequantised per row, fp64 accumulate. */
static void matvec_post(const orc_model *m, const orc_tensor *w, const float *x_in, float *y, int N, int K, float post) {
    float *xq = NULL; const float *x = x_in;
    if (m->act_quant && w->type != T_F32 && w->type != T_F16 && w->type != T_BF16) {
        xq = (float *)malloc((size_t)K * 4); quantise_act_256(x_in, K, xq); x = xq;
    }
    size_t rb = (size_t)K / type_block_elems(w->type) * type_block_bytes(w->type);
    #pragma omp parallel
    {
        float *row = (float *)malloc((size_t)K * 4);
        #pragma omp for schedule(static)
        for (int n = 0; n < N; n++) {
            orc_dequantize(w->type, w->data + (size_t)n * rb, K, row);
            double acc = 0; for (int k = 0; k < K; k++) acc += (double)row[k] * (double)x[k];
            y[n] = post == 1.0f ? (float)acc : (float)acc * post;
        }
        free(row);
    }
    free(xq);
}
static void matvec(const orc_model *m, const orc_tensor *w, const float *x_in, float *y, int N, int K) { matvec_post(m, w, x_in, y, N, K, 1.0f); }

static uint16_t fp32_to_fp16(float f);
static void round_f16(const float *x, float *y, int n) {
    for (int i = 0; i < n; i++) { float v = x[i]; if (v > 65504.f) v = 65504.f; if (v < -65504.f) v = -65504.f; y[i] = fp16_to_fp32(fp32_to_fp16(v)); }
}
/* RMSNorm.  Returns the factor the caller still has to apply to the results of the mat-vecs that consume y: 1 in the
 * plain mode.  In engine-numerics mode (act_quant) the engine quantises x*g block by block as the values arrive and
 * multiplies the mat-vec results by rsqrt(mean(x^2)+eps) afterwards (int8 block quantisation is invariant under a
 * common positive factor, so the codes are the ones x*sc*g would give up to fp32 rounding of the products); the
 * oracle follows the same order of operations. */
static float rmsnorm(const orc_model *m, const float *x, const float *g, float *y, int n, float eps) {
    double ss = 0; for (int i = 0; i < n; i++) ss += (double)x[i] * x[i];
    float sc = (float)(1.0 / sqrt(ss / n + (double)eps));
    if (m->act_quant) { for (int i = 0; i < n; i++) y[i] = x[i] * g[i]; return sc; }
    for (int i = 0; i < n; i++) y[i] = x[i] * sc * g[i];
    return 1.0f;
}

static const orc_tensor *layer_tensor(const orc_model *m, int l, const char *suffix) {
    char name[96]; snprintf(name, sizeof name, "blk.%d.%s", l, suffix); return find_tensor(m, name);
}

void orc_reset(orc_model *m) { m->n_past = 0; }

/* Feed one token at position n_past; writes logits[n_vocab] if logits != NULL.  Returns 0/-1. */
int orc_eval_token(orc_model *m, int32_t tok, float *logits) {
    const int D = m->n_embd, H = m->n_head, HK = m->n_head_kv, HD = m->head_dim, F = m->n_ff, V = m->n_vocab;
    const int KVD = HK * HD, pos = m->n_past;
    if (tok < 0 || tok >= V) return -1;
    if (pos >= m->kv_cap) {
        int ncap = m->kv_cap ? m->kv_cap * 2 : 256; size_t per = (size_t)m->n_layer * KVD;
        float *nk = (float *)calloc(per * ncap, 4), *nv = (float *)calloc(per * ncap, 4);
        for (int l = 0; l < m->n_layer && m->kv_cap; l++) {
            memcpy(nk + (size_t)l * ncap * KVD, m->kc + (size_t)l * m->kv_cap * KVD, (size_t)m->kv_cap * KVD * 4);
            memcpy(nv + (size_t)l * ncap * KVD, m->vc + (size_t)l * m->kv_cap * KVD, (size_t)m->kv_cap * KVD * 4);
        }
        free(m->kc); free(m->vc); m->kc = nk; m->vc = nv; m->kv_cap = ncap;
    }
    float *h = (float *)malloc((size_t)D * 4), *hx = (float *)malloc((size_t)D * 4), *x = (float *)malloc((size_t)D * 4), *q = (float *)malloc((size_t)H * HD * 4);
    float *att = (float *)malloc((size_t)H * HD * 4), *o = (float *)malloc((size_t)D * 4);
    float *g = (float *)malloc((size_t)F * 4), *u = (float *)malloc((size_t)F * 4), *nw = (float *)malloc((size_t)(D > F ? D : F) * 4);
    double *sc = (double *)malloc((size_t)(pos + 1) * 8);
    const orc_tensor *te = find_tensor(m, "token_embd.weight");
    size_t erb = (size_t)D / type_block_elems(te->type) * type_block_bytes(te->type);
    orc_dequantize(te->type, te->data + (size_t)tok * erb, D, h);
    if (m->emb_scale != 0.0f) for (int i = 0; i < D; i++) h[i] *= m->emb_scale;
    const orc_tensor *rf = find_tensor(m, "rope_freqs.weight");
    float kq_scale = m->attn_scale != 0.0f ? m->attn_scale : 1.0f / sqrtf((float)HD);
    float res = m->res_scale != 0.0f ? m->res_scale : 1.0f;
    int rc = 0;
    for (int l = 0; l < m->n_layer; l++) {
        const orc_tensor *an = layer_tensor(m, l, "attn_norm.weight"), *wq = layer_tensor(m, l, "attn_q.weight"),
            *wk = layer_tensor(m, l, "attn_k.weight"), *wv = layer_tensor(m, l, "attn_v.weight"), *wo = layer_tensor(m, l, "attn_output.weight"),
            *fn = layer_tensor(m, l, "ffn_norm.weight"), *wg = layer_tensor(m, l, "ffn_gate.weight"), *wu = layer_tensor(m, l, "ffn_up.weight"),
            *wd = layer_tensor(m, l, "ffn_down.weight");
        if (!an || !wq || !wk || !wv || !wo || !fn || !wg || !wu || !wd) { rc = -1; break; }
        float *kc = m->kc + ((size_t)l * m->kv_cap + pos) * KVD, *vc = m->vc + ((size_t)l * m->kv_cap + pos) * KVD;
        orc_dequantize(an->type, an->data, D, nw);
        float ps;
        if (m->xchg_f16) { round_f16(h, hx, D); ps = rmsnorm(m, hx, nw, x, D, m->rms_eps); } else ps = rmsnorm(m, h, nw, x, D, m->rms_eps);
        matvec_post(m, wq, x, q, H * HD, D, ps); matvec_post(m, wk, x, kc, KVD, D, ps); matvec_post(m, wv, x, vc, KVD, D, ps);
        /* RoPE, adjacent-pair ("NORM") form: theta_i = pos * base^(-2i/n_rot) / freq_factor_i */
        for (int hh = 0; hh < H + HK; hh++) {
            float *v = hh < H ? q + (size_t)hh * HD : kc + (size_t)(hh - H) * HD;
            for (int i = 0; i < m->n_rot / 2; i++) {
                double ff = 1.0; if (rf) { float t; orc_dequantize(rf->type, rf->data + (size_t)i * type_block_bytes(rf->type), 1, &t); ff = t; }
                double th = (double)pos * pow((double)m->rope_base, -2.0 * i / m->n_rot) / ff, c = cos(th), s = sin(th);
                float a = v[2 * i], b = v[2 * i + 1];
                v[2 * i] = (float)(a * c - b * s); v[2 * i + 1] = (float)(a * s + b * c);
            }
        }
        if (m->kv_f16) for (int i = 0; i < KVD; i++) { kc[i] = fp16_to_fp32(fp32_to_fp16(kc[i])); vc[i] = fp16_to_fp32(fp32_to_fp16(vc[i])); }
        for (int hh = 0; hh < H; hh++) {
            int kvh = hh / (H / HK); const float *qh = q + (size_t)hh * HD; double mx = -INFINITY;
            for (int p = 0; p <= pos; p++) {
                const float *kp = m->kc + ((size_t)l * m->kv_cap + p) * KVD + (size_t)kvh * HD;
                double s = 0; for (int d = 0; d < HD; d++) s += (double)qh[d] * kp[d];
                sc[p] = s * kq_scale; if (sc[p] > mx) mx = sc[p];
            }
            double sum = 0; for (int p = 0; p <= pos; p++) { sc[p] = exp(sc[p] - mx); sum += sc[p]; }
            for (int d = 0; d < HD; d++) {
                double a = 0;
                for (int p = 0; p <= pos; p++) a += sc[p] * m->vc[((size_t)l * m->kv_cap + p) * KVD + (size_t)kvh * HD + d];
                att[(size_t)hh * HD + d] = (float)(a / sum);
            }
        }
        if (m->xchg_f16) round_f16(att, att, H * HD);
        matvec(m, wo, att, o, D, H * HD);
        for (int i = 0; i < D; i++) h[i] += o[i] * res;
        orc_dequantize(fn->type, fn->data, D, nw);
        if (m->xchg_f16) { round_f16(h, hx, D); ps = rmsnorm(m, hx, nw, x, D, m->rms_eps); } else ps = rmsnorm(m, h, nw, x, D, m->rms_eps);
        matvec_post(m, wg, x, g, F, D, ps); matvec_post(m, wu, x, u, F, D, ps);
        for (int i = 0; i < F; i++) g[i] = (float)((double)g[i] / (1.0 + exp(-(double)g[i]))) * u[i];
        if (m->xchg_f16) round_f16(g, g, F);
        matvec(m, wd, g, o, D, F);
        for (int i = 0; i < D; i++) h[i] += o[i] * res;
    }
    m->n_past = pos + 1;
    if (!rc && logits) {
        const orc_tensor *on = find_tensor(m, "output_norm.weight"), *ow = find_tensor(m, "output.weight");
        if (!ow) ow = te;   /* tied embeddings (granite) */
        if (!on) rc = -1; else {
            orc_dequantize(on->type, on->data, D, nw);
            float ps;
            if (m->xchg_f16) { round_f16(h, hx, D); ps = rmsnorm(m, hx, nw, x, D, m->rms_eps); } else ps = rmsnorm(m, h, nw, x, D, m->rms_eps);
            matvec_post(m, ow, x, logits, V, D, ps);
            if (m->logit_scale != 0.0f) { float inv = 1.0f / m->logit_scale; for (int i = 0; i < V; i++) logits[i] *= inv; }
        }
    }
    free(h); free(hx); free(x); free(q); free(att); free(o); free(g); free(u); free(nw); free(sc);
    return rc;
}

/* Teacher-forced evaluation: feeds toks[0..n) from a fresh cache; logits_out is [n][V] if
 * all_logits else [V] (last position).  Returns 0/-1. */
int orc_forward(orc_model *m, const int32_t *toks, int n, float *logits_out, int all_logits) {
    orc_reset(m); int V = m->n_vocab;
    for (int i = 0; i < n; i++) {
        float *dst = all_logits ? logits_out + (size_t)i * V : (i == n - 1 ? logits_out : NULL);
        if (orc_eval_token(m, toks[i], dst)) return -1;
    }
    return 0;
}

/* Greedy generation: prompt then n_gen argmax tokens.  out_logits (optional) gets [n_gen][V]:
 * the logits each generated token was chosen from.  Lowest index wins ties. */
int orc_generate(orc_model *m, const int32_t *prompt, int n_prompt, int n_gen, int32_t *out, float *out_logits) {
    orc_reset(m); int V = m->n_vocab; float *lg = (float *)malloc((size_t)V * 4);
    for (int i = 0; i < n_prompt; i++) if (orc_eval_token(m, prompt[i], i == n_prompt - 1 ? lg : NULL)) { free(lg); return -1; }
    for (int t = 0; t < n_gen; t++) {
        int best = 0; for (int i = 1; i < V; i++) if (lg[i] > lg[best]) best = i;
        out[t] = best; if (out_logits) memcpy(out_logits + (size_t)t * V, lg, (size_t)V * 4);
        if (t + 1 < n_gen && orc_eval_token(m, best, lg)) { free(lg); return -1; }
    }
    free(lg); return 0;
}

/* Stand-alone mat-vec on raw ggml blocks: y[N] = W[N x K] . x[K]; act_quant as in orc_set_act_quant. */
int orc_matvec(int type, const void *raw, int N, int K, const float *x, float *y, int act_quant) {
    if (!type_block_elems(type) || K % type_block_elems(type)) return -1;
    orc_model m; memset(&m, 0, sizeof m); m.act_quant = act_quant;
    orc_tensor w; memset(&w, 0, sizeof w); w.type = type; w.data = (const uint8_t *)raw;
    matvec(&m, &w, x, y, N, K);
    return 0;
}
