// engine.cu -- model loading (GGUF / synthetic), context, and the decode path of one pipeline stage.
//
// Replaces, for the llama architecture, the reference's llm_load_tensors (llama.cpp:2684-2968), llama_new_context_with_model
// (:8707), llama_decode_internal_phased (:5461-5848: phase 0 = enqueue, phase 1 = results) and build_llama (:3872-3980).
// Kernel sequence per chunk of <= 16 tokens and per layer (fusion boundaries follow the reference's op order):
//   [x = down + ffn_inp] -> RMSNorm*w -> quantise          act_quant.cu   (K9 + K4 + K2)
//   Wq|Wk|Wv GEMV (one launch, rows concatenated)          gemv_*.cu      (K1)
//   RoPE(Q,K) -> fp16 Q, K/V cache store                   attention.cu   (K5 + K6)
//   masked multi-sequence attention                        attention.cu   (K7)
//   quantise -> Wo GEMV                                    (K2 + K1)
//   ffn_inp = o + x -> RMSNorm*w -> quantise               (K9 + K4 + K2)
//   Wgate|Wup GEMV -> silu(gate)*up -> quantise -> Wdown   (K1 + K8 + K2 + K1)
// then final RMSNorm -> lm_head GEMV -> argmax/top-8 on the device (K10).
#include "engine.cuh"
#include "gguf.cuh"

#include <math.h>
#include <string.h>
#include <sys/time.h>

#include <algorithm>

// kernels implemented in the other translation units
int  pib_gemv_internal(const pib_weight * w, const void * act_dev, int b, float * y_dev, int variant, const int * cancel, cudaStream_t st);
bool pib_gemv_pro_supported(const pib_weight * w, int b);
int  pib_gemv_pro(const pib_weight * w, int pro, const float * x, const float * x2, const float * addend, float * sum_out, int64_t xstride, float eps,
                  int b, float * y, const int * cancel, cudaStream_t st);
int  pib_silu_mul_quantize_act(int wtype, const float * gate_dev, const float * up_dev, int64_t row_stride, int b, int64_t K,
                               void * act_dev, const int * cancel, cudaStream_t st);
int  pib_quantize_act_c(int wtype, const float * x_dev, int b, int64_t K, void * act_dev, const int * cancel, cudaStream_t st);
int  pib_add_rmsnorm_quantize_act(int wtype, const float * x_dev, const float * addend_dev, float * sum_out_dev,
                                  const float * norm_w_dev, float eps, int b, int64_t K, void * act_dev, const int * cancel, cudaStream_t st);
int  pib_add_layernorm_quantize_act(int wtype, const float * x_dev, const float * addend_dev, const float * addend2_dev,
                                    float * sum_out_dev, const float * norm_w_dev, const float * norm_b_dev, float eps, int b,
                                    int64_t K, void * act_dev, const int * cancel, cudaStream_t st);
int  pib_gelu_quantize_act(int wtype, const float * x_dev, int b, int64_t K, void * act_dev, const int * cancel, cudaStream_t st);
void rope_table_build(std::vector<float2> & tab, int n_pos, int n_dims, int mode, float freq_base, float freq_scale);
bool pib_attention_supported(int hd, int n_head, int n_head_kv, int n_ctx);
int  pib_attention(const float * qkv, const float * vsep, const uint8_t * desc, const float2 * rope, int rope_n_pos, int mode, __half * kc, __half * vc,
                   const uint8_t * mask, const int32_t * scalars, int n_tokens, int n_head, int n_head_kv, int hd, int n_ctx, float * out,
                   const int * cancel, cudaStream_t st, bool cache_final);
int  pib_sample_topk(float * logits_dev, int n_rows, int n_vocab, const int32_t * bias_ids_dev, const float * bias_val_dev, int n_bias,
                     void * out_dev, const uint8_t * desc, const int * cancel, cudaStream_t st);
int  pib_gemm_q4k_tc(const pib_weight * w, const void * act_dev, int b, float * y_dev, void * xhat_dev, cudaStream_t st);
size_t pib_gemm_xhat_bytes(int64_t K, int b);
int  kv_device_mask(const KvDevice * kv, const int32_t * tpos_dev, const uint32_t * tokbits_dev, int n_tokens, uint8_t * mask_dev, cudaStream_t st);
int  pib_accept_walk(const void * rows_dev, const uint8_t * desc, void * out_dev, const int * cancel, cudaStream_t st);
pib_weight * pib_weight_alloc(int type, int64_t N, int64_t K);
void pib_weight_drop_planes(pib_weight * w);

// matrices the engine only multiplies with keep just their tiles (PIB_KEEP_PLANES=1: keep both forms, e.g. to run variant 1)
// exact mode (PIB_EXACT=1, read when a model is loaded and when a context is created): every matmul keeps the reference's
// summation order, so that logits are bit-identical to the CPU build's; the 32-element block types do that from the plane form
bool pib_exact_mode() { const char * e = getenv("PIB_EXACT"); return e && atoi(e) != 0; }

static void drop_planes_of(pib_model * m) {
    const char * keep = getenv("PIB_KEEP_PLANES");
    if (keep && atoi(keep) != 0) return;
    if (pib_exact_mode() && m->wtype != PIB_TYPE_Q4_K) return;
    // exact mode on a mixed file: Q4_K has its order-keeping tensor-core kernel, the other K-quants keep their planes for gemv_exact.cu
    auto drop = [&](pib_weight * w) { if (w && !(pib_exact_mode() && w->type != PIB_TYPE_Q4_K)) pib_weight_drop_planes(w); };
    drop(m->output);
    for (auto & L : m->layers) for (pib_weight * w : { L.wqkv, L.wv, L.wo, L.wgu, L.wdown }) drop(w);
}
int pib_weight_build_tiles(pib_weight * w);

static int64_t now_us() {
    struct timeval tv;
    gettimeofday(&tv, nullptr);
    return (int64_t) tv.tv_sec * 1000000 + tv.tv_usec;
}

// ---------------------------------------------------------------------------------------------------------------------
// model: GGUF
static float * upload_f32(const GgufTensor * t, int64_t n) {
    if (!t || t->type != 0 || t->ne[0] != n) return nullptr;
    float * d = nullptr;
    if (cudaMalloc(&d, (size_t) n * 4) != cudaSuccess) return nullptr;
    cudaMemcpy(d, t->data, (size_t) n * 4, cudaMemcpyHostToDevice);
    return d;
}

static size_t row_bytes_of(int type, int64_t K) {
    switch (type) {
        case PIB_TYPE_Q4_0: return (size_t) (K / 32) * 18;
        case PIB_TYPE_Q8_0: return (size_t) (K / 32) * 34;
        case PIB_TYPE_Q4_K: return (size_t) (K / 256) * 144;
        case PIB_TYPE_Q3_K: return (size_t) (K / 256) * 110;
        case PIB_TYPE_Q5_K: return (size_t) (K / 256) * 176;
        case PIB_TYPE_Q6_K: return (size_t) (K / 256) * 210;
    }
    return 0;
}

// upload the row-wise concatenation of several [N_i][K] tensors as one matrix
// (all of one type; `family` = the activation format the model multiplies with: a K-quant file may mix Q4_K and Q6_K)
static pib_weight * upload_concat(const std::vector<const GgufTensor *> & ts, int family, int64_t K) {
    int64_t N = 0;
    if (ts.empty() || !ts[0]) return nullptr;
    const int wtype = ts[0]->type;
    if (pib_act_family(wtype) != family) { pib_set_error("gguf: tensor type %d does not multiply with the activations of type %d", wtype, family); return nullptr; }
    for (auto t : ts) {
        if (!t || t->type != wtype || t->ne[0] != K) return nullptr;
        N += t->ne[1];
    }
    if (ts.size() == 1) return pib_weight_upload(wtype, ts[0]->data, N, K);
    const size_t rb = row_bytes_of(wtype, K);
    std::vector<uint8_t> tmp((size_t) N * rb);
    size_t off = 0;
    for (auto t : ts) { memcpy(tmp.data() + off, t->data, (size_t) t->ne[1] * rb); off += (size_t) t->ne[1] * rb; }
    return pib_weight_upload(wtype, tmp.data(), N, K);
}

pib_model * pib_model_load_gguf(const char * path, int layer_begin, int layer_end, bool want_embd, bool want_output) {
    GgufFile g;
    if (!g.open(path)) { pib_set_error("gguf: %s", g.error.c_str()); return nullptr; }
    const std::string arch = g.get_s("general.architecture", "");
    if (arch != "llama" && arch != "falcon") { pib_set_error("gguf: architecture '%s' not supported (llama, falcon)", arch.c_str()); return nullptr; }
    pib_model * m = new pib_model();
    m->arch = arch == "falcon" ? 1 : 0;
    m->name        = g.get_s("general.name", "");
    m->n_layer     = (int) g.get_i(arch + ".block_count", 0);
    m->n_embd      = (int) g.get_i(arch + ".embedding_length", 0);
    m->n_head      = (int) g.get_i(arch + ".attention.head_count", 0);
    m->n_head_kv   = (int) g.get_i(arch + ".attention.head_count_kv", m->n_head);
    m->n_ff        = (int) g.get_i(arch + ".feed_forward_length", 0);
    m->n_ctx_train = (int) g.get_i(arch + ".context_length", 2048);
    m->n_rot       = (int) g.get_i(arch + ".rope.dimension_count", m->n_head ? m->n_embd / m->n_head : 0);
    m->rms_eps        = (float) g.get_f(arch + ".attention.layer_norm_rms_epsilon", 1e-5);
    m->norm_eps       = (float) g.get_f(arch + ".attention.layer_norm_epsilon", 1e-5);
    m->rope_freq_base = (float) g.get_f(arch + ".rope.freq_base", 10000.0);
    if (g.kv.count("tokenizer.ggml.tokens")) m->tok_text = g.kv["tokenizer.ggml.tokens"].arr_s;
    // vocabulary kind and its default special ids (llama.cpp:2413-2468); an id outside the vocabulary falls back to the
    // default (:2505-2519)
    m->tok_model = g.get_s("tokenizer.ggml.model", "llama");
    const bool bpe = m->tok_model == "gpt2";
    if (bpe) {
        if (!g.kv.count("tokenizer.ggml.merges")) { pib_set_error("gguf: a gpt2 vocabulary without tokenizer.ggml.merges"); delete m; return nullptr; }
        m->tok_merges = g.kv["tokenizer.ggml.merges"].arr_s;
    }
    auto special_id = [&](const char * key, int def) {
        const int32_t id = (int32_t) g.get_i(key, def);                  // a u32 in the file, an int32 in the vocabulary
        return !m->tok_text.empty() && (size_t) ((int64_t) id + 1) > m->tok_text.size() ? def : (int) id;
    };
    m->tok_bos = special_id("tokenizer.ggml.bos_token_id", bpe ? 11 : 1);
    m->tok_eos = special_id("tokenizer.ggml.eos_token_id", bpe ? 11 : 2);
    { auto it = g.kv.find("tokenizer.ggml.add_bos_token"); if (it != g.kv.end() && it->second.type == 7) m->tok_add_bos = (int) it->second.u; }
    if (g.kv.count("tokenizer.ggml.scores")) for (double f : g.kv["tokenizer.ggml.scores"].arr_f) m->tok_score.push_back((float) f);
    if (g.kv.count("tokenizer.ggml.token_type")) for (int64_t i : g.kv["tokenizer.ggml.token_type"].arr_i) m->tok_type.push_back((int32_t) i);
    const GgufTensor * emb = g.tensor("token_embd.weight");
    // a vocabulary-only load (nothing resident) also accepts a file without tensors, like the reference's vocab_only
    const bool nothing_resident = !want_embd && !want_output && layer_begin == layer_end && layer_end >= 0;
    if ((!emb && !nothing_resident) || m->n_layer <= 0 || m->n_embd <= 0 || m->n_head <= 0) { pib_set_error("gguf: missing hyper-parameters / token_embd"); delete m; return nullptr; }
    m->n_vocab = emb ? (int) emb->ne[1] : (int) m->tok_text.size();
    // wtype = the activation family the model multiplies with; a Q3_K_M file (token_embd in Q3_K) is of the Q4_K / Q8_K family
    const int emb_type = emb ? emb->type : PIB_TYPE_Q8_0;
    m->wtype   = pib_act_family(emb_type);
    if (emb_type != PIB_TYPE_Q4_0 && emb_type != PIB_TYPE_Q8_0 && emb_type != PIB_TYPE_Q4_K && emb_type != PIB_TYPE_Q3_K && emb_type != PIB_TYPE_Q5_K &&
        emb_type != PIB_TYPE_Q6_K) {
        pib_set_error("gguf: token_embd of type %d not supported (Q4_0, Q8_0, Q3_K, Q4_K, Q5_K, Q6_K)", emb_type);
        delete m;
        return nullptr;
    }
    if (m->n_ff <= 0) { const GgufTensor * up = g.tensor("blk.0.ffn_up.weight"); if (up) m->n_ff = (int) up->ne[1]; }
    if (layer_end < 0 || layer_end > m->n_layer) layer_end = m->n_layer;        // < 0: all layers; begin == end: none (a head stage that only embeds)
    m->layer_begin = layer_begin;
    m->layer_end   = layer_end;
    m->layers.resize(m->n_layer);
    const int E = m->n_embd, FF = m->n_ff;
    bool ok = true;
    if (want_embd) { m->tok_embd = upload_concat({ emb }, m->wtype, E); ok = ok && m->tok_embd; }
    if (want_output) {
        m->output_norm = upload_f32(g.tensor("output_norm.weight"), E);
        m->output      = upload_concat({ g.tensor("output.weight") }, m->wtype, E);
        ok = ok && m->output_norm && m->output;
        if (m->arch == 1) { m->output_norm_b = upload_f32(g.tensor("output_norm.bias"), E); ok = ok && m->output_norm_b; }
    }
    for (int il = layer_begin; il < layer_end && ok; il++) {
        const std::string p = "blk." + std::to_string(il) + ".";
        pib_layer & L = m->layers[il];
        if (m->arch == 1) {                                   // falcon tensor table (reference llama.cpp:2769-2811)
            L.attn_norm   = upload_f32(g.tensor(p + "attn_norm.weight"), E);
            L.attn_norm_b = upload_f32(g.tensor(p + "attn_norm.bias"), E);
            if (g.tensor(p + "attn_norm_2.weight")) {
                L.attn_norm2   = upload_f32(g.tensor(p + "attn_norm_2.weight"), E);
                L.attn_norm2_b = upload_f32(g.tensor(p + "attn_norm_2.bias"), E);
                ok = ok && L.attn_norm2 && L.attn_norm2_b;
            }
            L.wqkv  = upload_concat({ g.tensor(p + "attn_qkv.weight") }, m->wtype, E);
            L.wo    = upload_concat({ g.tensor(p + "attn_output.weight") }, m->wtype, E);
            L.wgu   = upload_concat({ g.tensor(p + "ffn_up.weight") }, m->wtype, E);
            L.wdown = upload_concat({ g.tensor(p + "ffn_down.weight") }, m->wtype, FF);
            ok = ok && L.attn_norm && L.attn_norm_b && L.wqkv && L.wo && L.wgu && L.wdown;
            if (ok) m->weight_bytes += pib_weight_bytes(L.wqkv) + pib_weight_bytes(L.wo) + pib_weight_bytes(L.wgu) + pib_weight_bytes(L.wdown);
            continue;
        }
        L.attn_norm = upload_f32(g.tensor(p + "attn_norm.weight"), E);
        L.ffn_norm  = upload_f32(g.tensor(p + "ffn_norm.weight"), E);
        {
            const GgufTensor * tq = g.tensor(p + "attn_q.weight"), * tk = g.tensor(p + "attn_k.weight"), * tv = g.tensor(p + "attn_v.weight");
            if (tq && tk && tv && tv->type != tq->type) {       // Q4_K_M: attn_v in Q6_K (or Q5_K) next to Q4_K attn_q / attn_k
                L.wqkv = upload_concat({ tq, tk }, m->wtype, E);
                L.wv   = upload_concat({ tv }, m->wtype, E);
                ok = ok && L.wv;
                if (L.wv) m->weight_bytes += pib_weight_bytes(L.wv);
            } else L.wqkv = upload_concat({ tq, tk, tv }, m->wtype, E);
        }
        L.wo    = upload_concat({ g.tensor(p + "attn_output.weight") }, m->wtype, E);
        L.wgu   = upload_concat({ g.tensor(p + "ffn_gate.weight"), g.tensor(p + "ffn_up.weight") }, m->wtype, E);
        L.wdown = upload_concat({ g.tensor(p + "ffn_down.weight") }, m->wtype, FF);
        ok = L.attn_norm && L.ffn_norm && L.wqkv && L.wo && L.wgu && L.wdown;
        if (ok) m->weight_bytes += pib_weight_bytes(L.wqkv) + pib_weight_bytes(L.wo) + pib_weight_bytes(L.wgu) + pib_weight_bytes(L.wdown);
    }
    if (!ok) {
        if (!*pib_last_error()) pib_set_error("gguf: missing or mistyped tensor");
        pib_model_free(m);
        return nullptr;
    }
    if (m->output) m->weight_bytes += pib_weight_bytes(m->output);
    drop_planes_of(m);
    return m;
}

// ---------------------------------------------------------------------------------------------------------------------
// model: synthetic (random init generated directly as quantised blocks on the device)
__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ uint32_t hash3(uint32_t seed, uint32_t a, uint32_t b) { return hash32(seed ^ hash32(a * 0x9E3779B9U + hash32(b + 0x85ebca6bU))); }

// one thread per 16 bytes of the qs plane / per scale entry; `rowmap` (optional) gives the logical row whose random
// stream a physical row uses (token_embd row t and output row pi(t) share one stream in the bigram flavour)
__global__ void synth_fill_kernel(int type, uint8_t * qs, uint16_t * d, int64_t N, int64_t K, uint32_t seed, const int32_t * rowmap,
                                  float dscale) {
    const int64_t qs_row = type == PIB_TYPE_Q8_0 ? K : K / 2;
    const int64_t nq     = N * qs_row / 16;
    const int64_t i      = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nq) {
        const int64_t  row = i / (qs_row / 16), c = i % (qs_row / 16);
        const uint32_t lr  = rowmap ? (uint32_t) rowmap[row] : (uint32_t) row;
        uint4 v;
        v.x = hash3(seed, lr, (uint32_t) (c * 4 + 0)); v.y = hash3(seed, lr, (uint32_t) (c * 4 + 1));
        v.z = hash3(seed, lr, (uint32_t) (c * 4 + 2)); v.w = hash3(seed, lr, (uint32_t) (c * 4 + 3));
        reinterpret_cast<uint4 *>(qs)[i] = v;
    }
    const int64_t nd = type == PIB_TYPE_Q4_K ? N * (K / 256) : N * (K / 32);
    if (i < nd) {
        const int64_t  bpr = type == PIB_TYPE_Q4_K ? K / 256 : K / 32;
        const int64_t  row = i / bpr, c = i % bpr;
        const uint32_t lr  = rowmap ? (uint32_t) rowmap[row] : (uint32_t) row;
        const uint32_t h   = hash3(seed ^ 0xA5A5A5A5u, lr, (uint32_t) c);
        const float    u   = 0.5f + (float) (h & 0xffff) / 65536.0f;       // U(0.5, 1.5)
        if (type != PIB_TYPE_Q4_K) {
            // Q8_0: value = d * q, q ~ U{-128..127} (std 74);  Q4_0: d * (q - 8) (std 4.6)
            const float dd = type == PIB_TYPE_Q8_0 ? dscale * u / 74.0f : dscale * u / 4.6f;
            d[i] = __half_as_ushort(__float2half_rn(dd));
        } else {
            // value = d*sc*q - dmin*m with m = sc and dmin = 7.5*d  =>  d*sc*(q - 7.5): zero mean, std = d*sc*4.6
            uint8_t * hdr = reinterpret_cast<uint8_t *>(d) + i * 16;
            const float dd = dscale * u / (32.0f * 4.6f);
            const uint16_t hd = __half_as_ushort(__float2half_rn(dd)), hm = __half_as_ushort(__float2half_rn(7.5f * __half2float(__float2half_rn(dd))));
            uint32_t s[8];
            for (int j = 0; j < 8; j++) s[j] = 16u + (hash3(seed ^ 0x5C5C5C5Cu, lr, (uint32_t) (c * 8 + j)) & 31u);   // 16..47
            // 12 packed bytes such that get_scale_min_k4 yields sc_j = m_j = s[j] (reference ggml-quants.c:1446-1454)
            uint32_t q[12];
            for (int j = 0; j < 4; j++) { q[j] = s[j] | ((s[j + 4] >> 4) << 6); q[j + 4] = s[j] | ((s[j + 4] >> 4) << 6); }
            for (int j = 4; j < 8; j++) q[j + 4] = (s[j] & 0xFu) | ((s[j] & 0xFu) << 4);
            uint4 w;
            w.x = (uint32_t) hd | ((uint32_t) hm << 16);
            w.y = q[0] | (q[1] << 8) | (q[2] << 16) | (q[3] << 24);
            w.z = q[4] | (q[5] << 8) | (q[6] << 16) | (q[7] << 24);
            w.w = q[8] | (q[9] << 8) | (q[10] << 16) | (q[11] << 24);
            *reinterpret_cast<uint4 *>(hdr) = w;
        }
    }
}

static pib_weight * synth_weight(int type, int64_t N, int64_t K, uint32_t seed, const int32_t * rowmap_dev, float dscale) {
    pib_weight * w = pib_weight_alloc(type, N, K);
    if (!w) return nullptr;
    const int64_t qs_row = type == PIB_TYPE_Q8_0 ? K : K / 2;
    const int64_t n      = N * qs_row / 16;
    synth_fill_kernel<<<(unsigned) ((n + 255) / 256), 256>>>(type, w->qs, w->d, N, K, seed, rowmap_dev, dscale);
    if (cudaGetLastError() != cudaSuccess || pib_weight_build_tiles(w) != PIB_OK) { pib_weight_free(w); return nullptr; }
    return w;
}

static float * const_f32(int64_t n, float v) {
    std::vector<float> h((size_t) n, v);
    float * d = nullptr;
    if (cudaMalloc(&d, (size_t) n * 4) != cudaSuccess) return nullptr;
    cudaMemcpy(d, h.data(), (size_t) n * 4, cudaMemcpyHostToDevice);
    return d;
}

// flavour 0: iid weights.  flavour 1: "bigram": token_embd[t] = g_t and output[pi(t)] = g_t, layer weights small, so
// the model predicts pi(last token); a draft built with agree < 1 uses pi on a fraction `agree` of the tokens and a
// cyclic shift of pi on the rest, giving a known per-token acceptance probability.
pib_model * pib_model_synthetic(int n_layer, int n_embd, int n_head, int n_head_kv, int n_ff, int n_vocab, int wtype, uint64_t seed,
                                int flavour, float agree, int layer_begin, int layer_end, bool want_embd, bool want_output) {
    const int blk = wtype == PIB_TYPE_Q4_K ? 256 : 32;
    if (n_embd % blk || n_ff % blk || n_embd % n_head || n_head % n_head_kv) { pib_set_error("synthetic model: bad shape"); return nullptr; }
    // flavour: low nibble 0 iid / 1 bigram; 0x10 = falcon architecture, 0x20 = with the second attention norm (40B/180B layout)
    const bool falcon = (flavour & 0x10) != 0, norm2 = (flavour & 0x20) != 0;
    flavour &= 0xF;
    pib_model * m = new pib_model();
    m->arch = falcon ? 1 : 0;
    m->name = "synthetic";
    m->n_layer = n_layer; m->n_embd = n_embd; m->n_head = n_head; m->n_head_kv = n_head_kv; m->n_ff = n_ff; m->n_vocab = n_vocab;
    m->n_rot = n_embd / n_head; m->n_ctx_train = 4096; m->wtype = wtype;
    if (layer_end < 0 || layer_end > n_layer) layer_end = n_layer;              // < 0: all layers; begin == end: none
    m->layer_begin = layer_begin; m->layer_end = layer_end;
    m->layers.resize(n_layer);
    const int E = n_embd, GQA = m->n_embd_gqa(), FF = n_ff;
    // permutations are a function of the vocabulary size only (shared by target and draft)
    std::vector<int32_t> pi(n_vocab), inv(n_vocab);
    {
        uint64_t s = 0x9E3779B97F4A7C15ull ^ (uint64_t) n_vocab;
        auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; };
        for (int i = 0; i < n_vocab; i++) pi[i] = i;
        for (int i = n_vocab - 1; i > 0; i--) { int j = (int) (rnd() % (uint64_t) (i + 1)); std::swap(pi[i], pi[j]); }
        if (agree < 1.0f) {
            std::vector<int> dis;
            uint64_t s2 = 0xD1B54A32D192ED03ull ^ (uint64_t) n_vocab;
            auto rnd2 = [&]() { s2 ^= s2 << 13; s2 ^= s2 >> 7; s2 ^= s2 << 17; return s2; };
            for (int t = 0; t < n_vocab; t++) if ((double) (rnd2() % 1000000) / 1e6 >= agree) dis.push_back(t);
            if (dis.size() > 1) {
                const int32_t first = pi[dis[0]];
                for (size_t k = 0; k + 1 < dis.size(); k++) pi[dis[k]] = pi[dis[k + 1]];
                pi[dis.back()] = first;
            }
        }
        for (int t = 0; t < n_vocab; t++) inv[pi[t]] = t;
    }
    int32_t * inv_dev = nullptr;
    cudaMalloc(&inv_dev, (size_t) n_vocab * 4);
    cudaMemcpy(inv_dev, inv.data(), (size_t) n_vocab * 4, cudaMemcpyHostToDevice);
    const uint32_t s32   = (uint32_t) (seed * 2654435761u);
    // layer weight std: iid flavour 0.02; bigram flavour 0.1/sqrt(K) so that a layer perturbs the residual stream by ~1%
    // of its norm whatever the width (80 layers of a 70B-shaped model then still predict pi(last token))
    const float    lstdE = flavour == 1 ? 0.1f / sqrtf((float) E) : 0.02f, lstdF = flavour == 1 ? 0.1f / sqrtf((float) FF) : 0.02f;
    const uint32_t embseed = flavour == 1 ? 0xE3B0C442u : s32 ^ 0x11111111u;   // bigram: the same g_t for target and draft
    bool ok = true;
    if (want_embd) { m->tok_embd = synth_weight(wtype, n_vocab, E, embseed, nullptr, flavour == 1 ? 1.0f : 0.02f); ok = ok && m->tok_embd; }
    if (want_output) {
        m->output_norm = const_f32(E, 1.0f);
        if (falcon) { m->output_norm_b = const_f32(E, 0.0f); ok = ok && m->output_norm_b; }
        m->output = flavour == 1 ? synth_weight(wtype, n_vocab, E, embseed, inv_dev, 1.0f) : synth_weight(wtype, n_vocab, E, s32 ^ 0x22222222u, nullptr, 0.02f);
        ok = ok && m->output && m->output_norm;
    }
    for (int il = layer_begin; il < layer_end && ok; il++) {
        pib_layer & L = m->layers[il];
        const uint32_t ls = s32 ^ (uint32_t) (il * 7919 + 13);
        L.attn_norm = const_f32(E, 1.0f);
        if (falcon) {
            L.attn_norm_b = const_f32(E, 0.0f);
            if (norm2) { L.attn_norm2 = const_f32(E, 1.0f); L.attn_norm2_b = const_f32(E, 0.0f); }
        } else L.ffn_norm = const_f32(E, 1.0f);
        L.wqkv  = synth_weight(wtype, E + 2 * GQA, E, ls ^ 0x1u, nullptr, lstdE);
        L.wo    = synth_weight(wtype, E, E, ls ^ 0x2u, nullptr, lstdE);
        L.wgu   = synth_weight(wtype, (falcon ? 1 : 2) * FF, E, ls ^ 0x3u, nullptr, lstdE);
        L.wdown = synth_weight(wtype, E, FF, ls ^ 0x4u, nullptr, lstdF);
        ok = L.attn_norm && (falcon ? (L.attn_norm_b && (!norm2 || (L.attn_norm2 && L.attn_norm2_b))) : L.ffn_norm != nullptr) && L.wqkv && L.wo && L.wgu && L.wdown;
        if (ok) m->weight_bytes += pib_weight_bytes(L.wqkv) + pib_weight_bytes(L.wo) + pib_weight_bytes(L.wgu) + pib_weight_bytes(L.wdown);
    }
    cudaDeviceSynchronize();
    cudaFree(inv_dev);
    if (!ok || cudaGetLastError() != cudaSuccess) {
        if (!*pib_last_error()) pib_set_error("synthetic model: allocation failed");
        pib_model_free(m);
        return nullptr;
    }
    if (m->output) m->weight_bytes += pib_weight_bytes(m->output);
    drop_planes_of(m);
    return m;
}

void pib_model_free(pib_model * m) {
    if (!m) return;
    pib_weight_free(m->tok_embd);
    pib_weight_free(m->output);
    if (m->output_norm) cudaFree(m->output_norm);
    if (m->output_norm_b) cudaFree(m->output_norm_b);
    for (auto & L : m->layers) {
        for (float * f : { L.attn_norm_b, L.attn_norm2, L.attn_norm2_b }) if (f) cudaFree(f);
        if (L.attn_norm) cudaFree(L.attn_norm);
        if (L.ffn_norm) cudaFree(L.ffn_norm);
        pib_weight_free(L.wqkv); pib_weight_free(L.wv); pib_weight_free(L.wo); pib_weight_free(L.wgu); pib_weight_free(L.wdown);
    }
    delete m;
}

// ---------------------------------------------------------------------------------------------------------------------
// context
#define CK(expr) do { if ((expr) != cudaSuccess) { pib_set_error("%s:%d %s: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(cudaGetLastError())); goto fail; } } while (0)

pib_context * pib_context_new(pib_model * m, uint32_t n_ctx, int max_seq, pib_stage stage) {
    if (!pib_attention_supported(m->head_dim(), m->n_head, m->n_head_kv, (int) n_ctx)) {
        pib_set_error("context: head_dim %d / n_head %d / n_head_kv %d / n_ctx %u not supported by the attention kernel",
                      m->head_dim(), m->n_head, m->n_head_kv, n_ctx);
        return nullptr;
    }
    pib_context * c = new pib_context();
    c->model = m; c->n_ctx = n_ctx; c->max_seq = max_seq < 1 ? 1 : max_seq; c->n_words = (c->max_seq + 31) / 32; c->stage = stage;
    c->t_start_us = now_us();
    const int    E = m->n_embd, GQA = m->n_embd_gqa(), FF = m->n_ff, B = PIB_CHUNK;
    const int    nl = m->layer_end - m->layer_begin;
    const size_t kvb = (size_t) nl * n_ctx * GQA * sizeof(__half);
    int lo, hi;
    std::vector<float2> tab;
    size_t actb;
    CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CK(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, lo));
    CK(cudaStreamCreateWithPriority(&c->ctrl, cudaStreamNonBlocking, hi));
    CK(cudaEventCreateWithFlags(&c->cancel_ev, cudaEventDisableTiming));
    c->kv.init(n_ctx, c->max_seq);
    if (kv_device_init(&c->kvd, n_ctx, c->n_words) != PIB_OK) goto fail;
    CK(cudaMalloc(&c->kcache, kvb ? kvb : 16));
    CK(cudaMalloc(&c->vcache, kvb ? kvb : 16));
    CK(cudaMemset(c->kcache, 0, kvb));
    CK(cudaMemset(c->vcache, 0, kvb));
    CK(cudaMalloc(&c->x, (size_t) B * E * 4));
    CK(cudaMalloc(&c->x2, (size_t) B * E * 4));
    CK(cudaMalloc(&c->qkv, (size_t) B * (E + 2 * GQA) * 4));
    CK(cudaMalloc(&c->vsep, (size_t) B * GQA * 4));
    CK(cudaMalloc(&c->att, (size_t) B * E * 4));
    CK(cudaMalloc(&c->proj, (size_t) B * E * 4));
    CK(cudaMalloc(&c->gu, (size_t) B * 2 * FF * 4));
    CK(cudaMalloc(&c->mask, (size_t) B * n_ctx));
    actb = std::max(pib_act_bytes(m->wtype, E, B), pib_act_bytes(m->wtype, FF, B));
    CK(cudaMalloc(&c->act, actb));
    c->rope_n_pos = (int) std::max<uint32_t>(n_ctx, (uint32_t) m->n_ctx_train) + 64;
    rope_table_build(tab, c->rope_n_pos, m->n_rot, m->arch == 1 ? 2 : 0, m->rope_freq_base, m->rope_freq_scale);    // falcon: NeoX pairs
    CK(cudaMalloc(&c->rope, tab.size() * sizeof(float2)));
    CK(cudaMemcpy(c->rope, tab.data(), tab.size() * sizeof(float2), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&c->cancel_table, 128 * sizeof(int)));          // [0] current run slot, [1..64] flags, [100] the constant 1
    CK(cudaMemset(c->cancel_table, 0, 128 * sizeof(int)));
    c->cancel_one = c->cancel_table + 100;
    { int one = 1; CK(cudaMemcpy(c->cancel_one, &one, 4, cudaMemcpyHostToDevice)); }
    CK(cudaMallocHost(&c->run_id_host, 64 * sizeof(int)));
    for (int i = 0; i < 64; i++) c->run_id_host[i] = i;
    c->desc_stride = pib_desc_stride(c->n_words);
    // ring of 256 slots for multi-chunk decodes; a single-chunk decode stages its descriptor in host slot 257 + run slot and
    // always lands in device slot 256: one fixed address, ordered by the stream, so that its CUDA graph serves every run
    c->desc_slots  = 256;
    CK(cudaMallocHost(&c->desc_host, c->desc_stride * (c->desc_slots + 65)));
    CK(cudaMalloc(&c->desc_dev, c->desc_stride * (c->desc_slots + 65)));
    c->desc_ev.assign(c->desc_slots + 65, nullptr);
    { const char * g = getenv("PIB_GRAPHS"); c->use_graphs = !g || atoi(g) != 0; }
    {
        // prologue-fused GEMVs (PIB_FUSE_PRO=1; off by default): measured on the 1.1B draft, 204 -> 160 launches per step buy 1.6 % at one
        // token (0.754 -> 0.742 ms) and lose at two (0.769 -> 0.864): the step is a chain of dependent memory round trips, and moving a
        // quantiser into the matmul moves its round trips along with it
        const char * g = getenv("PIB_FUSE_PRO");
        c->fuse_pro = g ? atoi(g) : 0;
    }
    { const char * g = getenv("PIB_PROMPT_GEMM"); c->prompt_gemm = (g && atoi(g) == 0) || pib_exact_mode() ? 0 : 1; }
    c->ops_cap = 1 << 16;
    CK(cudaMallocHost(&c->ops_host, sizeof(KvOp) * c->ops_cap));
    CK(cudaMalloc(&c->ops_dev, sizeof(KvOp) * c->ops_cap));
    c->runs.resize(64);
    for (int i = 0; i < (int) c->runs.size(); i++) c->runs[i].id = i;
    if (stage.size > 1 && pib_pipe_hop_setup(c) != PIB_OK) goto fail;
    return c;
fail:
    pib_context_free(c);
    return nullptr;
}

void pib_context_free(pib_context * c) {
    if (!c) return;
    if (c->stream) cudaStreamSynchronize(c->stream);
    pib_pipe_hop_free(c);
    for (void * b : { (void *) c->pb.x, (void *) c->pb.x2, (void *) c->pb.qkv, (void *) c->pb.att, (void *) c->pb.proj, (void *) c->pb.gu,
                      (void *) c->pb.mask, c->pb.act, c->pb.xhat }) if (b) cudaFree(b);
    kv_device_free(&c->kvd);
    void * bufs[] = { c->vsep, c->kcache, c->vcache, c->x, c->x2, c->qkv, c->att, c->proj, c->gu, c->mask, c->act, c->rope,
                      c->cancel_table, c->desc_dev, c->ops_dev, c->bias_ids_dev, c->bias_val_dev, c->logits_fetch_dev };
    for (void * b : bufs) if (b) cudaFree(b);
    if (c->desc_host) cudaFreeHost(c->desc_host);
    if (c->run_id_host) cudaFreeHost(c->run_id_host);
    if (c->ops_host) cudaFreeHost(c->ops_host);
    for (auto & r : c->runs) {
        if (r.done) cudaEventDestroy(r.done);
        if (r.logits) cudaFree(r.logits);
        if (r.res_dev) cudaFree(r.res_dev);
        if (r.res_host) cudaFreeHost(r.res_host);
        if (r.acc_dev) cudaFree(r.acc_dev);
        if (r.acc_host) cudaFreeHost(r.acc_host);
    }
    for (auto e : c->desc_ev) if (e) cudaEventDestroy(e);
    if (c->cancel_ev) cudaEventDestroy(c->cancel_ev);
    for (auto & g : c->graphs) if (g.second.exec) cudaGraphExecDestroy(g.second.exec);
    if (c->stream) cudaStreamDestroy(c->stream);
    if (c->ctrl) cudaStreamDestroy(c->ctrl);
    delete c;
}

// ---------------------------------------------------------------------------------------------------------------------
// KV op-log front end: see kvlog_* in kv_meta.cu (lazy, coalescing; shared with the CPU-only test hooks)
void pib_kv_push(pib_context * c, const KvOp & op) { kvlog_push(c->pending, op); }
void pib_kv_flush_host(pib_context * c) { kvlog_flush(c->kv, c->pending, c->devops); }
static std::vector<KvOp> & devq(pib_context * c) { return c->devops; }

// ---------------------------------------------------------------------------------------------------------------------
// decode
static inline int dev_slot(const pib_context * c, int host_slot) { return host_slot > c->desc_slots ? c->desc_slots : host_slot; }

static int take_desc_slot(pib_context * c, const pib_run * run, int n_chunks) {
    int s;
    if (run && n_chunks == 1) return c->desc_slots + 1 + run->id;  // host staging of the run slot (free: its last run has finished)
    else { s = c->desc_next; c->desc_next = (c->desc_next + 1) % c->desc_slots; }
    if (c->desc_ev[s]) cudaEventSynchronize(c->desc_ev[s]);
    else cudaEventCreateWithFlags(&c->desc_ev[s], cudaEventDisableTiming);
    return s;
}

__global__ void add_kernel(const float * __restrict__ a, const float * __restrict__ b, float * __restrict__ o, int64_t n,
                           const int * __restrict__ cancel) {
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) o[i] = __fadd_rn(a[i], b[i]);
}

// falcon: l_out = (ffn_out + attn_out) + layer input, in the reference's order (llama.cpp:4197-4201); o may alias c
__global__ void add3_kernel(const float * a, const float * b, const float * c, float * o, int64_t n, const int * __restrict__ cancel) {
    pib_pdl_wait();
    pib_pdl_trigger();
    if (PIB_CANCELLED(cancel)) return;
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) o[i] = __fadd_rn(__fadd_rn(a[i], b[i]), c[i]);
}

// hooks installed by pipeline.cu on multi-GPU runs (nullptr on a single GPU)
int (*pib_hop_recv)(pib_context * c, float * x, int n_tokens, const int * cancel) = nullptr;
int (*pib_hop_send)(pib_context * c, const float * x, int n_tokens, const int * cancel) = nullptr;

// debugging aid (PIB_DUMP_DIR=<dir>, graphs off): every intermediate of the llama graph goes to <dir>/<name>.<layer>.bin as float32,
// under the names the oracle uses for the same tensors (PO_DUMP_DIR, oracle/pib_oracle.c), so the two can be compared step by step
static void dump_dev(const pib_context * c, const char * name, int il, const float * dev, size_t rows, size_t cols, size_t ld) {
    static const char * dir = getenv("PIB_DUMP_DIR");
    if (!dir) return;
    cudaStreamSynchronize(c->stream);
    std::vector<float> h(rows * cols);
    cudaMemcpy2D(h.data(), cols * 4, dev, ld * 4, cols * 4, rows, cudaMemcpyDeviceToHost);
    char path[512];
    snprintf(path, sizeof(path), "%s/%s.%d.bin", dir, name, il);
    if (FILE * f = fopen(path, "wb")) { fwrite(h.data(), 4, h.size(), f); fclose(f); }
}

// `part`: 1 = embedding + layers (everything whose addresses are the same for every run: graph material), 2 = the last
// stage's tail (final norm, lm_head, sampling: writes the run's own logits / result rows), 3 = both
static int run_chunk_body(pib_context * c, pib_run & run, int chunk, int ntok, int slot, const int * cancel, int part) {
    pib_model *    m    = c->model;
    const int      E = m->n_embd, GQA = m->n_embd_gqa(), FF = m->n_ff, hd = m->head_dim(), wt = m->wtype, var = c->gemv_variant;
    cudaStream_t   st   = c->stream;
    const uint8_t * desc = c->desc_dev + (size_t) dev_slot(c, slot) * c->desc_stride;
    const int32_t * tok  = reinterpret_cast<const int32_t *>(desc + PIB_DESC_TOK_OFF);
    const int32_t * tpos = reinterpret_cast<const int32_t *>(desc + PIB_DESC_POS_OFF);
    const uint32_t * tbits = reinterpret_cast<const uint32_t *>(desc + PIB_DESC_BITS_OFF);
    int rc;
#define RC(e) do { rc = (e); if (rc != PIB_OK) return rc; c->n_kernels++; } while (0)
    // matmul launch, optionally bracketed by events (kind: 0 qkv, 1 o, 2 gate|up, 3 down, 4 lm_head)
#define GEMV(kind, il_, w, out) do {                                                                                   \
        const bool pr__ = c->prof_now && (il_) == c->prof_layer && c->prof.size() < 8192;                                             \
        pib_context::ProfRec r__ = { (kind), ntok, nullptr, nullptr };                                                 \
        if (pr__) { cudaEventCreate(&r__.e0); cudaEventCreate(&r__.e1); cudaEventRecord(r__.e0, st); }                 \
        if (!(skip_mask >> (kind) & 1)) RC(pib_gemv_internal((w), c->act, ntok, (out), var, cancel, st));              \
        if (ntok > 8 && var != 2) c->n_kernels++;                                                                      \
        if (pr__) { cudaEventRecord(r__.e1, st); c->prof.push_back(r__); }                                             \
    } while (0)
    // any other launch of the profiled layer, bracketed the same way (kind: 5 attn-norm+quant, 6 rope+kv store, 7 attention,
    // 8 quant of the attention output, 9 residual+ffn-norm+quant, 10 silu*up+quant)
#define TL(kind, il_, call) do {                                                                                       \
        const bool pr__ = c->prof_now && (il_) == c->prof_layer && c->prof.size() < 8192;                                             \
        pib_context::ProfRec r__ = { (kind), ntok, nullptr, nullptr };                                                 \
        if (pr__) { cudaEventCreate(&r__.e0); cudaEventCreate(&r__.e1); cudaEventRecord(r__.e0, st); }                 \
        if (!(skip_mask >> (kind) & 1)) RC(call);                                                                      \
        if (pr__) { cudaEventRecord(r__.e1, st); c->prof.push_back(r__); }                                             \
    } while (0)
    // timing experiments only (tools/pass_breakdown.py): leave out every launch of the kinds in the mask; the results are garbage
    static const unsigned skip_mask = getenv("PIB_DEBUG_SKIP_KINDS") ? (unsigned) strtoul(getenv("PIB_DEBUG_SKIP_KINDS"), nullptr, 0) : 0u;
    if (part & 1) {
        if (c->hop_peer && !c->stage.first()) RC(pib_hop_recv_launch(c, ntok, st));
        RC(kv_device_mask(&c->kvd, tpos, tbits, ntok, c->mask, st));
        if (c->stage.first()) RC(pib_get_rows(m->tok_embd, tok, ntok, c->x, st));
    }
    if (m->arch == 1) {
        // ---- falcon (reference build_falcon, llama.cpp:4106-4225) ----------------------------------------------------------
        //   [x = (ffn_out + attn_out) + x_prev] -> LayerNorm*w+b -> quantise      (attn_norm_2 when the layer has one, else attn_norm)
        //   Wqkv GEMV; [LayerNorm(attn_norm) -> quantise;] Wup GEMV               the MLP reads the attention norm's output
        //   NeoX RoPE, K/V store, attention, quantise, Wo GEMV -> attn_out
        //   gelu(up) -> quantise -> Wdown GEMV -> ffn_out
        // The layer input lives in `cur`; the attention output and then the next layer's input go to `other` (ping-pong), so
        // that no kernel reads a buffer another CTA of the same launch is rewriting.
        const int nl = m->layer_end - m->layer_begin;
        float *   cur = c->x, * other = c->att;
        bool      pend = false;
        if (!(part & 1)) { pend = nl > 0; if (nl > 0 && ((nl - 1) & 1)) std::swap(cur, other); }     // the state the layer loop leaves behind
        for (int il = m->layer_begin; (part & 1) && il < m->layer_end; il++) {
            const pib_layer & L  = m->layers[il];
            const size_t      lo = (size_t) (il - m->layer_begin) * c->n_ctx * GQA;
            const float *     nw = L.attn_norm2 ? L.attn_norm2 : L.attn_norm, * nb = L.attn_norm2 ? L.attn_norm2_b : L.attn_norm_b;
            if (pend) { TL(5, il, pib_add_layernorm_quantize_act(wt, c->x2, c->proj, cur, other, nw, nb, m->norm_eps, ntok, E, c->act, cancel, st)); std::swap(cur, other); }
            else      { TL(5, il, pib_add_layernorm_quantize_act(wt, cur, nullptr, nullptr, nullptr, nw, nb, m->norm_eps, ntok, E, c->act, cancel, st)); }
            GEMV(0, il, L.wqkv, c->qkv);
            if (L.attn_norm2) TL(9, il, pib_add_layernorm_quantize_act(wt, cur, nullptr, nullptr, nullptr, L.attn_norm, L.attn_norm_b, m->norm_eps, ntok, E, c->act, cancel, st));
            GEMV(2, il, L.wgu, c->gu);
            TL(7, il, pib_attention(c->qkv, nullptr, desc, c->rope, c->rope_n_pos, 2, c->kcache + lo, c->vcache + lo, c->mask, c->kvd.scalars, ntok, m->n_head,
                                    m->n_head_kv, hd, (int) c->n_ctx, other, cancel, st, true));
            TL(8, il, pib_quantize_act_c(wt, other, ntok, E, c->act, cancel, st));
            GEMV(1, il, L.wo, c->proj);
            TL(10, il, pib_gelu_quantize_act(wt, c->gu, ntok, FF, c->act, cancel, st));
            GEMV(3, il, L.wdown, c->x2);
            pend = true;
        }
        if (c->stage.last() && !(part & 2)) return PIB_OK;
        if (!c->stage.last() && !(part & 1)) return PIB_OK;
        if (c->stage.last()) {
            if (pend) RC(pib_add_layernorm_quantize_act(wt, c->x2, c->proj, cur, other, m->output_norm, m->output_norm_b, m->norm_eps, ntok, E, c->act, cancel, st));
            else      RC(pib_add_layernorm_quantize_act(wt, cur, nullptr, nullptr, nullptr, m->output_norm, m->output_norm_b, m->norm_eps, ntok, E, c->act, cancel, st));
            float * lg = run.logits + (size_t) chunk * PIB_CHUNK * m->n_vocab;
            GEMV(4, c->prof_layer >= 0 ? c->prof_layer : -2, m->output, lg);
            RC(pib_sample_topk(lg, ntok, m->n_vocab, c->bias_ids_dev, c->bias_val_dev, c->n_bias, run.res_dev + (size_t) chunk * PIB_CHUNK,
                               desc, cancel, st));
            if (run.has_walk) RC(pib_accept_walk(run.res_dev, desc, run.acc_dev, cancel, st));
        } else if (c->hop_peer) {
            if (pend) RC(pib_hop_send_launch(c, c->x2, c->proj, cur, ntok, st));
            else      RC(pib_hop_send_launch(c, cur, nullptr, nullptr, ntok, st));
        } else if (pend) {
            const int64_t n = (int64_t) ntok * E;
            pib_launch(add3_kernel, dim3((unsigned) ((n + 255) / 256)), dim3(256), 0, st, c->x2, c->proj, cur, c->x, n, cancel);
            c->n_kernels++;
        }
        return PIB_OK;
    }
    // x holds layer_inp; `pending_add` = a down-projection result still to be added to x2 (ffn_inp) to form the next x
    bool pending_add = (part & 1) ? false : m->layer_end > m->layer_begin;
    for (int il = m->layer_begin; (part & 1) && il < m->layer_end; il++) {
        const pib_layer & L  = m->layers[il];
        const size_t      lo = (size_t) (il - m->layer_begin) * c->n_ctx * GQA;
        // PIB_FUSE_PRO=1: with one or two tokens the QKV and attention-output matmuls form their own quantised activations
        // (gemv_q4k_pro_kernel: prologue 1 = [add +] RMSNorm, 0 = plain), 9 -> 7 launches per layer.  gate|up (704 CTAs on the draft, each
        // redoing the 2048-wide norm) and down (silu over 5632 values per CTA) lose outright -- 17 against 8.9 and 8.0 us -- and are never fused.
        const bool fuse = c->fuse_pro && var == 0 && !L.wv && ntok <= 2 && pib_gemv_pro_supported(L.wqkv, ntok) && pib_gemv_pro_supported(L.wo, ntok) &&
                          L.wqkv->N <= 16 * 592 && L.wo->N <= 16 * 592;
#define GEMVP(kind, ...) do {                                                                                          \
        const bool pr__ = c->prof_now && il == c->prof_layer && c->prof.size() < 8192;                                 \
        pib_context::ProfRec r__ = { (kind), ntok, nullptr, nullptr };                                                 \
        if (pr__) { cudaEventCreate(&r__.e0); cudaEventCreate(&r__.e1); cudaEventRecord(r__.e0, st); }                 \
        if (!(skip_mask >> (kind) & 1)) RC(pib_gemv_pro(__VA_ARGS__, cancel, st));                                     \
        if (pr__) { cudaEventRecord(r__.e1, st); c->prof.push_back(r__); }                                             \
    } while (0)
        if (fuse) {
            if (pending_add) GEMVP(0, L.wqkv, 1, c->proj, L.attn_norm, c->x2, c->x, 0, m->rms_eps, ntok, c->qkv);
            else             GEMVP(0, L.wqkv, 1, c->x, L.attn_norm, nullptr, nullptr, 0, m->rms_eps, ntok, c->qkv);
            TL(7, il, pib_attention(c->qkv, nullptr, desc, c->rope, c->rope_n_pos, 0, c->kcache + lo, c->vcache + lo, c->mask, c->kvd.scalars, ntok,
                                    m->n_head, m->n_head_kv, hd, (int) c->n_ctx, c->att, cancel, st, true));
            GEMVP(1, L.wo, 0, c->att, nullptr, nullptr, nullptr, 0, 0.f, ntok, c->proj);
            TL(9, il, pib_add_rmsnorm_quantize_act(wt, c->proj, c->x, c->x2, L.ffn_norm, m->rms_eps, ntok, E, c->act, cancel, st));   // x2 = ffn_inp
            GEMV(2, il, L.wgu, c->gu);
            TL(10, il, pib_silu_mul_quantize_act(wt, c->gu, c->gu + FF, 2 * (int64_t) FF, ntok, FF, c->act, cancel, st));
            GEMV(3, il, L.wdown, c->proj);
            pending_add = true;
            continue;
        }
#undef GEMVP
        if (pending_add) TL(5, il, pib_add_rmsnorm_quantize_act(wt, c->proj, c->x2, c->x, L.attn_norm, m->rms_eps, ntok, E, c->act, cancel, st));
        else             TL(5, il, pib_add_rmsnorm_quantize_act(wt, c->x, nullptr, nullptr, L.attn_norm, m->rms_eps, ntok, E, c->act, cancel, st));
        GEMV(0, il, L.wqkv, c->qkv);
        if (L.wv) GEMV(0, il, L.wv, c->vsep);
        dump_dev(c, "inp", il, c->x, ntok, E, E);
        dump_dev(c, "q", il, c->qkv, ntok, E, E + (L.wv ? 1 : 2) * GQA);
        dump_dev(c, "k", il, c->qkv + E, ntok, GQA, E + (L.wv ? 1 : 2) * GQA);
        if (!L.wv) dump_dev(c, "v", il, c->qkv + E + GQA, ntok, GQA, E + 2 * GQA);
        TL(7, il, pib_attention(c->qkv, L.wv ? c->vsep : nullptr, desc, c->rope, c->rope_n_pos, 0, c->kcache + lo, c->vcache + lo, c->mask, c->kvd.scalars, ntok,
                                m->n_head, m->n_head_kv, hd, (int) c->n_ctx, c->att, cancel, st, true));
        dump_dev(c, "att", il, c->att, ntok, E, E);
        TL(8, il, pib_quantize_act_c(wt, c->att, ntok, E, c->act, cancel, st));
        GEMV(1, il, L.wo, c->proj);
        dump_dev(c, "o", il, c->proj, ntok, E, E);
        TL(9, il, pib_add_rmsnorm_quantize_act(wt, c->proj, c->x, c->x2, L.ffn_norm, m->rms_eps, ntok, E, c->act, cancel, st));   // x2 = ffn_inp
        dump_dev(c, "ffi", il, c->x2, ntok, E, E);
        GEMV(2, il, L.wgu, c->gu);
        dump_dev(c, "gate", il, c->gu, ntok, FF, 2 * (size_t) FF);
        dump_dev(c, "up", il, c->gu + FF, ntok, FF, 2 * (size_t) FF);
        TL(10, il, pib_silu_mul_quantize_act(wt, c->gu, c->gu + FF, 2 * (int64_t) FF, ntok, FF, c->act, cancel, st));
        GEMV(3, il, L.wdown, c->proj);
        dump_dev(c, "down", il, c->proj, ntok, E, E);
        pending_add = true;
    }
    if (c->stage.last() && !(part & 2)) return PIB_OK;
    if (!c->stage.last() && !(part & 1)) return PIB_OK;
    if (c->stage.last()) {
        if (pending_add) RC(pib_add_rmsnorm_quantize_act(wt, c->proj, c->x2, c->x, m->output_norm, m->rms_eps, ntok, E, c->act, cancel, st));
        else             RC(pib_add_rmsnorm_quantize_act(wt, c->x, nullptr, nullptr, m->output_norm, m->rms_eps, ntok, E, c->act, cancel, st));
        float * lg = run.logits + (size_t) chunk * PIB_CHUNK * m->n_vocab;
        GEMV(4, c->prof_layer >= 0 ? c->prof_layer : -2, m->output, lg);
        RC(pib_sample_topk(lg, ntok, m->n_vocab, c->bias_ids_dev, c->bias_val_dev, c->n_bias, run.res_dev + (size_t) chunk * PIB_CHUNK,
                           desc, cancel, st));
        if (run.has_walk) RC(pib_accept_walk(run.res_dev, desc, run.acc_dev, cancel, st));
    } else if (c->hop_peer) {
        // the stage's output goes straight into the next stage's input slot: residual add fused with the store over NVLink
        if (pending_add) RC(pib_hop_send_launch(c, c->proj, c->x2, nullptr, ntok, st));
        else             RC(pib_hop_send_launch(c, c->x, nullptr, nullptr, ntok, st));
    } else {
        if (pending_add) {
            const int64_t n = (int64_t) ntok * E;
            pib_launch(add_kernel, dim3((unsigned) ((n + 255) / 256)), dim3(256), 0, st, c->proj, c->x2, c->x, n, cancel);
            c->n_kernels++;
        }
    }
#undef TL
#undef GEMV
#undef RC
    return PIB_OK;
}

// One chunk on this stage: receive the activations (stages > 0), run the layers, send them on (stages < last).  The
// layer part of a single-chunk decode is replayed from a CUDA graph captured on the second use of its n_tokens: every
// address a kernel of that part gets is the same for every run (one device descriptor slot, the cancel block, the work
// buffers) and everything that varies between decodes lives in the descriptor, which is device memory.  One graph launch replaces
// ~10 launches per layer on the host, which is what bounds the small draft model.
static int run_chunk(pib_context * c, pib_run & run, int chunk, int ntok, int slot) {
    const int * cancel = c->cancel_table;            // block address (see PIB_CANCELLED); the current-run cell was set by enqueue_decode
    int rc;
    if (!c->hop_peer && !c->stage.first() && pib_hop_recv && (rc = pib_hop_recv(c, c->x, ntok, cancel)) != PIB_OK) return rc;
    // per-launch event timing (bench.py roofline) samples one decode in eight, launched directly; the others keep their graphs
    c->prof_now = c->prof_layer >= 0 && (c->n_decode & 7) == 0 && slot > c->desc_slots;      // verification batches only (single chunk), not prompt chunks
    const bool graphable = c->use_graphs && !c->prof_now && slot > c->desc_slots;           // single-chunk decode: fixed device slot
    bool done = false;
    if (graphable) {
        pib_context::GraphEntry & g = c->graphs[ntok];
        g.uses++;
        if (!g.exec && !g.failed && g.uses >= 2) {
            const uint64_t k0 = c->n_kernels;
            cudaGraph_t graph = nullptr;
            bool ok = cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeRelaxed) == cudaSuccess;
            if (ok) {
                rc = run_chunk_body(c, run, chunk, ntok, slot, cancel, 1);
                ok = cudaStreamEndCapture(c->stream, &graph) == cudaSuccess && rc == PIB_OK && graph != nullptr;
            }
            if (ok) ok = cudaGraphInstantiate(&g.exec, graph, 0) == cudaSuccess;
            if (graph) cudaGraphDestroy(graph);
            g.n_kernels = (int) (c->n_kernels - k0);
            c->n_kernels = k0;
            if (!ok) { cudaGetLastError(); g.exec = nullptr; g.failed = true; }
        }
        if (g.exec) {
            if (cudaGraphLaunch(g.exec, c->stream) != cudaSuccess) { pib_set_error("decode: graph launch failed: %s", cudaGetErrorString(cudaGetLastError())); return PIB_ERR_CUDA; }
            c->n_kernels += (uint64_t) g.n_kernels;
            c->n_graph_launches++;
            if ((rc = run_chunk_body(c, run, chunk, ntok, slot, cancel, 2)) != PIB_OK) return rc;
            done = true;
        }
    }
    if (!done) {
        // a decode whose launches are timed with events runs in plain stream order: with PDL a kernel starts (and prefetches
        // weights) before the event in front of it fires, and the event pair would under-report its duration
        if (c->prof_now) pib_pdl_suspend(true);
        rc = run_chunk_body(c, run, chunk, ntok, slot, cancel, 3);
        if (c->prof_now) pib_pdl_suspend(false);
        if (rc != PIB_OK) return rc;
    }
    if (!c->hop_peer && !c->stage.last() && pib_hop_send && (rc = pib_hop_send(c, c->x, ntok, cancel)) != PIB_OK) return rc;
    if (c->hop_peer && !c->stage.last()) pib_pipe_hop_account((size_t) ntok * c->model->n_embd * 4);
    return PIB_OK;
}

static int enqueue_decode(pib_context * c, pib_run * run, int n_tokens, int batch_id, const std::vector<int> & slots);

// ---- wide batches (prompt ingestion): the matmuls of up to PIB_PROMPT_MAX tokens at once on the tcgen05 GEMM -----------------------
// The reference takes a prompt of up to n_batch = 512 tokens in one ggml_compute_forward_mul_mat per matrix (common.h:52, ggml.c:9388).
// Here the per-token kernels (norm / quantise, RoPE + KV store, attention, ...) keep their 16-token descriptors, but every matmul of a
// layer runs once for the whole pass.  Used for llama-architecture models whose matrices are all Q4_K, for decodes of more than two
// chunks, outside exact mode (the GEMM is tolerance-checked, not bit-exact: gemm_q4k_tc.cu).
static bool prompt_gemm_ok(const pib_context * c, int n_tokens) {
    const pib_model * m = c->model;
    if (!c->prompt_gemm || c->gemv_variant != 0 || m->arch != 0 || m->wtype != PIB_TYPE_Q4_K || n_tokens <= 2 * PIB_CHUNK) return false;
    for (int il = m->layer_begin; il < m->layer_end; il++) {
        const pib_layer & L = m->layers[il];
        if (L.wv) return false;
        for (const pib_weight * w : { L.wqkv, L.wo, L.wgu, L.wdown }) if (!w || w->type != PIB_TYPE_Q4_K || !w->tiles) return false;
    }
    if (c->stage.last() && (!m->output || m->output->type != PIB_TYPE_Q4_K || !m->output->tiles)) return false;
    return true;
}

static int prompt_bufs(pib_context * c) {
    if (c->pb.ready) return PIB_OK;
    const pib_model * m = c->model;
    const size_t P = PIB_PROMPT_MAX, E = m->n_embd, GQA = m->n_embd_gqa(), FF = m->n_ff, KM = std::max(E, FF);
    auto & b = c->pb;
    if (cudaMalloc(&b.x, P * E * 4) != cudaSuccess || cudaMalloc(&b.x2, P * E * 4) != cudaSuccess || cudaMalloc(&b.qkv, P * (E + 2 * GQA) * 4) != cudaSuccess ||
        cudaMalloc(&b.att, P * E * 4) != cudaSuccess || cudaMalloc(&b.proj, P * E * 4) != cudaSuccess || cudaMalloc(&b.gu, P * 2 * FF * 4) != cudaSuccess ||
        cudaMalloc(&b.mask, P * c->n_ctx) != cudaSuccess ||
        cudaMalloc(&b.act, pib_act_bytes(PIB_TYPE_Q4_K, (int64_t) KM, (int) P)) != cudaSuccess || cudaMalloc(&b.xhat, pib_gemm_xhat_bytes((int64_t) KM, (int) P)) != cudaSuccess) {
        cudaGetLastError();
        pib_set_error("decode: out of memory for the prompt path's work buffers");
        return PIB_ERR_NOMEM;
    }
    b.ready = true;
    return PIB_OK;
}

// one pass of `ntok` tokens = chunks [ch0, ch0 + nch) of the run
static int run_prompt_pass(pib_context * c, pib_run & run, int ch0, int nch, int ntok, const std::vector<int> & slots) {
    pib_model *  m  = c->model;
    const int    E = m->n_embd, GQA = m->n_embd_gqa(), FF = m->n_ff, hd = m->head_dim(), wt = m->wtype;
    cudaStream_t st = c->stream;
    const int *  cancel = c->cancel_table;
    auto & B = c->pb;
    int rc;
#define RC(e) do { rc = (e); if (rc != PIB_OK) return rc; c->n_kernels++; } while (0)
    auto desc_of = [&](int k) { return c->desc_dev + (size_t) dev_slot(c, slots[ch0 + k]) * c->desc_stride; };
    auto nt_of   = [&](int k) { return std::min(PIB_CHUNK, ntok - k * PIB_CHUNK); };
    for (int k = 0; k < nch; k++) {
        const uint8_t * desc = desc_of(k);
        const size_t    off  = (size_t) k * PIB_CHUNK;
        if (c->stage.first()) RC(pib_get_rows(m->tok_embd, reinterpret_cast<const int32_t *>(desc + PIB_DESC_TOK_OFF), nt_of(k), B.x + off * E, st));
        else if (c->hop_peer) { RC(pib_hop_recv_launch(c, nt_of(k), st)); cudaMemcpyAsync(B.x + off * E, c->x, (size_t) nt_of(k) * E * 4, cudaMemcpyDeviceToDevice, st); }
        else if (pib_hop_recv) RC(pib_hop_recv(c, B.x + off * E, nt_of(k), cancel));
        RC(kv_device_mask(&c->kvd, reinterpret_cast<const int32_t *>(desc + PIB_DESC_POS_OFF), reinterpret_cast<const uint32_t *>(desc + PIB_DESC_BITS_OFF), nt_of(k),
                          B.mask + off * c->n_ctx, st));
    }
    bool pending_add = false;
    for (int il = m->layer_begin; il < m->layer_end; il++) {
        const pib_layer & L  = m->layers[il];
        const size_t      lo = (size_t) (il - m->layer_begin) * c->n_ctx * GQA;
        if (pending_add) RC(pib_add_rmsnorm_quantize_act(wt, B.proj, B.x2, B.x, L.attn_norm, m->rms_eps, ntok, E, B.act, cancel, st));
        else             RC(pib_add_rmsnorm_quantize_act(wt, B.x, nullptr, nullptr, L.attn_norm, m->rms_eps, ntok, E, B.act, cancel, st));
        RC(pib_gemm_q4k_tc(L.wqkv, B.act, ntok, B.qkv, B.xhat, st));
        for (int k = 0; k < nch; k++) {          // chunk k + 1 reads chunk k's cells from the cache: stream order
            const size_t off = (size_t) k * PIB_CHUNK;
            RC(pib_attention(B.qkv + off * (E + 2 * GQA), nullptr, desc_of(k), c->rope, c->rope_n_pos, 0, c->kcache + lo, c->vcache + lo, B.mask + off * c->n_ctx,
                             c->kvd.scalars, nt_of(k), m->n_head, m->n_head_kv, hd, (int) c->n_ctx, B.att + off * E, cancel, st, k == 0));
        }
        RC(pib_quantize_act_c(wt, B.att, ntok, E, B.act, cancel, st));
        RC(pib_gemm_q4k_tc(L.wo, B.act, ntok, B.proj, B.xhat, st));
        RC(pib_add_rmsnorm_quantize_act(wt, B.proj, B.x, B.x2, L.ffn_norm, m->rms_eps, ntok, E, B.act, cancel, st));
        RC(pib_gemm_q4k_tc(L.wgu, B.act, ntok, B.gu, B.xhat, st));
        RC(pib_silu_mul_quantize_act(wt, B.gu, B.gu + FF, 2 * (int64_t) FF, ntok, FF, B.act, cancel, st));
        RC(pib_gemm_q4k_tc(L.wdown, B.act, ntok, B.proj, B.xhat, st));
        pending_add = true;
    }
    if (c->stage.last()) {
        if (pending_add) RC(pib_add_rmsnorm_quantize_act(wt, B.proj, B.x2, B.x, m->output_norm, m->rms_eps, ntok, E, B.act, cancel, st));
        else             RC(pib_add_rmsnorm_quantize_act(wt, B.x, nullptr, nullptr, m->output_norm, m->rms_eps, ntok, E, B.act, cancel, st));
        float * lg = run.logits + (size_t) ch0 * PIB_CHUNK * m->n_vocab;
        RC(pib_gemm_q4k_tc(m->output, B.act, ntok, lg, B.xhat, st));
        RC(pib_sample_topk(lg, ntok, m->n_vocab, c->bias_ids_dev, c->bias_val_dev, c->n_bias, run.res_dev + (size_t) ch0 * PIB_CHUNK, desc_of(0), cancel, st));
    } else {
        // the stage's output, chunk by chunk, to the next stage (its hop slots hold one 16-token chunk each)
        for (int k = 0; k < nch; k++) {
            const size_t off = (size_t) k * PIB_CHUNK * E;
            if (c->hop_peer) {
                if (pending_add) RC(pib_hop_send_launch(c, B.proj + off, B.x2 + off, nullptr, nt_of(k), st));
                else             RC(pib_hop_send_launch(c, B.x + off, nullptr, nullptr, nt_of(k), st));
                pib_pipe_hop_account((size_t) nt_of(k) * E * 4);
            } else if (pib_hop_send) {
                if (pending_add) {
                    const int64_t n = (int64_t) nt_of(k) * E;
                    pib_launch(add_kernel, dim3((unsigned) ((n + 255) / 256)), dim3(256), 0, st, B.proj + off, B.x2 + off, B.x + off, n, cancel);
                    c->n_kernels++;
                }
                RC(pib_hop_send(c, B.x + off, nt_of(k), cancel));
            }
        }
    }
#undef RC
    return PIB_OK;
}

static pib_run * take_run(pib_context * c, int n_tokens) {
    for (auto & r : c->runs) {
        if (r.in_use || r.id == c->last_finished) continue;
        if (!r.done) cudaEventCreateWithFlags(&r.done, cudaEventDisableTiming);
        if (!r.acc_dev && (cudaMalloc(&r.acc_dev, sizeof(pib_accept_result)) != cudaSuccess || cudaMallocHost(&r.acc_host, sizeof(pib_accept_result)) != cudaSuccess)) {
            pib_set_error("decode: out of memory for a run slot");
            cudaGetLastError();
            return nullptr;
        }
        const int cap = ((n_tokens + PIB_CHUNK - 1) / PIB_CHUNK) * PIB_CHUNK;
        if ((c->stage.last() || c->stage.first()) && r.cap_tokens < cap) {
            if (r.res_dev) { cudaStreamSynchronize(c->stream); cudaFree(r.logits); cudaFree(r.res_dev); cudaFreeHost(r.res_host); }
            r.logits = nullptr; r.res_dev = nullptr; r.res_host = nullptr;
            if ((c->stage.last() && cudaMalloc(&r.logits, (size_t) cap * c->model->n_vocab * 4) != cudaSuccess) ||
                cudaMalloc(&r.res_dev, (size_t) cap * sizeof(pib_row_result)) != cudaSuccess ||
                cudaMallocHost(&r.res_host, (size_t) cap * sizeof(pib_row_result)) != cudaSuccess) {
                pib_set_error("decode: out of memory for logits of %d tokens", cap);
                cudaGetLastError();
                return nullptr;
            }
            r.cap_tokens = cap;
        }
        return &r;
    }
    pib_set_error("decode: more than %zu runs in flight", c->runs.size());
    return nullptr;
}

static bool slow_trace() { static int on = -1; if (on < 0) on = getenv("PIB_TRACE_SLOW") ? 1 : 0; return on != 0; }
#define SLOW_MARK(tag) do { if (slow_trace()) { const int64_t n__ = now_us(); if (n__ - t_mark > 500) fprintf(stderr, "[slow] rank-stage %d batch %d: %s took %lld us\n", c->stage.rank, batch_id, tag, (long long) (n__ - t_mark)); t_mark = n__; } } while (0)

int pib_decode_start(pib_context * c, int n_tokens, const int32_t * token, const int32_t * pos, const int32_t * n_seq_id,
                     const int32_t * const * seq_id, int batch_id) {
    if (n_tokens <= 0) { pib_set_error("decode: n_tokens == 0"); return -1; }
    int64_t t_mark = slow_trace() ? now_us() : 0;
    pib_model * m = c->model;
    pib_kv_flush_host(c);
    SLOW_MARK("kv flush (host mirror)");
    std::vector<KvOp> & dq = devq(c);
    const int W = c->n_words;
    // per-token sequence bitsets (the reference's SEQ_IDS message, ggml-mpi.c:318-334, as rows of n_words words)
    std::vector<uint32_t> tb((size_t) n_tokens * W, 0u);
    for (int i = 0; i < n_tokens; i++)
        for (int j = 0; j < n_seq_id[i]; j++) {
            const int s = seq_id[i][j];
            if (s < 0 || s >= W * 32) { pib_set_error("decode: sequence id %d out of range (max_seq %d)", s, c->max_seq); return -1; }
            tb[(size_t) i * W + (s >> 5)] |= 1u << (s & 31);
        }
    for (int i = 0; i < n_tokens; i++)
        if (c->stage.first() && (token[i] < 0 || token[i] >= m->n_vocab)) { pib_set_error("decode: token id %d out of range", token[i]); return -1; }
    // llama.cpp:5548-5571: head-reset heuristic, find_slot, attended range n
    const bool found = kvlog_decode_begin(c->kv, dq, n_tokens, pos, tb.data());
    SLOW_MARK("find slot");
    pib_run *  run   = found ? take_run(c, n_tokens) : nullptr;
    SLOW_MARK("take run");
    if (!found) pib_set_error("decode: no KV slot for %d tokens", n_tokens);
    if (!run) return -1;      // bookkeeping done so far (flushed ops, head moves) stays queued for the device
    const uint32_t head = c->kv.head;
    const int n_chunks = (n_tokens + PIB_CHUNK - 1) / PIB_CHUNK;
    std::vector<int> slots(n_chunks);
    for (int ch = 0; ch < n_chunks; ch++) {
        const int t0 = ch * PIB_CHUNK, nt = std::min(PIB_CHUNK, n_tokens - t0);
        const int s  = take_desc_slot(c, run, n_chunks);
        slots[ch]    = s;
        uint8_t * d  = c->desc_host + (size_t) s * c->desc_stride;
        int32_t * hdr = reinterpret_cast<int32_t *>(d + PIB_DESC_HDR_OFF);
        memset(d, 0, c->desc_stride);
        hdr[PIB_DH_NTOK] = nt; hdr[PIB_DH_KVHEAD] = (int32_t) (head + t0); hdr[PIB_DH_BATCHID] = batch_id; hdr[PIB_DH_NKV] = (int32_t) c->kv.n;
        memcpy(d + PIB_DESC_TOK_OFF, token + t0, sizeof(int32_t) * nt);
        memcpy(d + PIB_DESC_POS_OFF, pos + t0, sizeof(int32_t) * nt);
        memcpy(d + PIB_DESC_BITS_OFF, tb.data() + (size_t) t0 * W, sizeof(uint32_t) * (size_t) nt * W);
        if (ch == 0 && n_chunks == 1 && c->tree_pending.n_branches > 0) {           // the acceptance table rides in the descriptor
            const pib_accept_table & T = c->tree_pending;
            hdr[PIB_DH_TREE_NP] = T.n_branches; hdr[PIB_DH_TREE_KEEP] = T.keep; hdr[PIB_DH_TREE_SPEC] = T.speculative;
            int32_t * tl = reinterpret_cast<int32_t *>(d + PIB_DESC_TREE_OFF);
            memcpy(tl, T.len, sizeof(T.len));
            memcpy(tl + PIB_TREE_NP, T.rows, sizeof(T.rows));
            memcpy(tl + PIB_TREE_NP + PIB_TREE_NP * PIB_TREE_DEPTH, T.toks, sizeof(T.toks));
        }
        dq.push_back({ KVOP_ALLOC, (int32_t) (head + t0), nt, (int32_t) c->kv.n, s, 0 });
        cudaMemcpyAsync(c->desc_dev + (size_t) dev_slot(c, s) * c->desc_stride, d, c->desc_stride, cudaMemcpyHostToDevice, c->stream);
        c->h2d_bytes += c->desc_stride;
    }
    dq.push_back({ KVOP_ADVANCE, n_tokens, (int32_t) head, 0, 0, 0 });
    kvlog_decode_end(c->kv, n_tokens);                                    // llama.cpp:5708-5713
    SLOW_MARK("descriptors");
    if (c->stage.size > 1 && c->stage.first()) {          // tell the other stages first, so that their receives are posted early
        std::vector<const uint8_t *> dp(n_chunks);
        for (int ch = 0; ch < n_chunks; ch++) dp[ch] = c->desc_host + (size_t) slots[ch] * c->desc_stride;
        if (pib_pipe_ship_decode(batch_id, n_tokens, n_chunks, dp.data(), c->desc_stride, slots.data(), dq.data(), (int) dq.size()) != PIB_OK) return -1;
    }
    SLOW_MARK("ship record");
    run->has_walk = n_chunks == 1 && c->tree_pending.n_branches > 0;
    c->tree_pending.n_branches = 0;
    const int rid = enqueue_decode(c, run, n_tokens, batch_id, slots);
    SLOW_MARK("enqueue");
    return rid;
}

// ship the queued device-level op-log to the GPU and replay it on the KV replica
int pib_apply_devops(pib_context * c) {
    std::vector<KvOp> & dq = devq(c);
    const int n = (int) dq.size();
    if (n == 0) return PIB_OK;
    if (n > c->ops_cap) { pib_set_error("decode: KV op-log overflow (%d ops)", n); return PIB_ERR_ARG; }
    if (c->ops_next + n > c->ops_cap) { cudaStreamSynchronize(c->stream); c->ops_next = 0; }
    memcpy(c->ops_host + c->ops_next, dq.data(), sizeof(KvOp) * n);
    for (int i = 0; i < n; i++)                               // the log names host staging slots; the kernel reads device slots
        if (c->ops_host[c->ops_next + i].op == KVOP_ALLOC) c->ops_host[c->ops_next + i].p1 = dev_slot(c, c->ops_host[c->ops_next + i].p1);
    cudaMemcpyAsync(c->ops_dev + c->ops_next, c->ops_host + c->ops_next, sizeof(KvOp) * n, cudaMemcpyHostToDevice, c->stream);
    c->h2d_bytes += sizeof(KvOp) * n;
    if (kv_device_apply(&c->kvd, c->ops_dev + c->ops_next, n, c->desc_dev, c->desc_stride, c->stream) != PIB_OK) return PIB_ERR_CUDA;
    c->n_kernels++;
    c->ops_next += n;
    dq.clear();
    return PIB_OK;
}

// head: bring every replica (its own GPU's and the other stages') up to date with the KV ops issued so far, without a decode
int pib_kv_sync_replicas(pib_context * c) {
    pib_kv_flush_host(c);
    std::vector<KvOp> & dq = devq(c);
    if (dq.empty()) return PIB_OK;
    if (c->stage.size > 1 && c->stage.first() &&
        pib_pipe_ship_decode(0, 0, 0, nullptr, c->desc_stride, nullptr, dq.data(), (int) dq.size()) != PIB_OK) return PIB_ERR_NOMEM;
    return pib_apply_devops(c);
}

// device part of a decode, common to the head and the other stages: replay the op-log on the KV replica, run the chunks,
// hand the sampled rows to whoever needs them
static int enqueue_decode(pib_context * c, pib_run * run, int n_tokens, int batch_id, const std::vector<int> & slots) {
    const int n_chunks = (int) slots.size();
    int64_t t_mark = slow_trace() ? now_us() : 0;
    if (pib_apply_devops(c) != PIB_OK) return -1;
    SLOW_MARK("  apply ops");
    run->in_use = true; run->batch_id = batch_id; run->n_tokens = n_tokens; run->failed = false;
    cudaMemcpyAsync(c->cancel_table, c->run_id_host + run->id, sizeof(int), cudaMemcpyHostToDevice, c->stream);   // "this run executes now"
    if (prompt_gemm_ok(c, n_tokens) && prompt_bufs(c) == PIB_OK) {
        const int per = PIB_PROMPT_MAX / PIB_CHUNK;
        for (int ch0 = 0; ch0 < n_chunks && !run->failed; ch0 += per) {
            const int nch = std::min(per, n_chunks - ch0), ntok = std::min(PIB_PROMPT_MAX, n_tokens - ch0 * PIB_CHUNK);
            if (run_prompt_pass(c, *run, ch0, nch, ntok, slots) != PIB_OK) run->failed = true;
        }
        for (int ch = 0; ch < n_chunks; ch++) if (c->desc_ev[slots[ch]]) cudaEventRecord(c->desc_ev[slots[ch]], c->stream);
    } else
    for (int ch = 0; ch < n_chunks; ch++) {
        const int nt = std::min(PIB_CHUNK, n_tokens - ch * PIB_CHUNK);
        if (run_chunk(c, *run, ch, nt, slots[ch]) != PIB_OK) { run->failed = true; break; }
        if (c->desc_ev[slots[ch]]) cudaEventRecord(c->desc_ev[slots[ch]], c->stream);
    }
    SLOW_MARK("  chunks");
    if (c->stage.last() && !run->failed) {
        if (c->stage.size > 1) {
            if (pib_pipe_send_results(c, *run) != PIB_OK) run->failed = true;
        } else {
            // with an acceptance table only the walk's result crosses to the host; the rows stay on the device (fetched on demand)
            if (run->has_walk) {
                cudaMemcpyAsync(run->acc_host, run->acc_dev, sizeof(pib_accept_result), cudaMemcpyDeviceToHost, c->stream);
                c->d2h_bytes += sizeof(pib_accept_result);
                run->rows_on_host = false;
            } else {
                cudaMemcpyAsync(run->res_host, run->res_dev, sizeof(pib_row_result) * (size_t) n_tokens, cudaMemcpyDeviceToHost, c->stream);
                c->d2h_bytes += sizeof(pib_row_result) * (size_t) n_tokens;
                run->rows_on_host = true;
            }
        }
    }
    cudaEventRecord(run->done, c->stream);
    if (c->stage.size > 1 && c->stage.first() && !run->failed) {
        // the rows come back from the last stage on the result stream; `done` then means "rows are in pinned host memory"
        if (pib_pipe_post_result_recv(c, *run) != PIB_OK) run->failed = true;
    }
    SLOW_MARK("  result recv post");
    c->inflight.push_back(run->id);
    c->n_decode++;
    return run->id;
}

int pib_decode_replay(pib_context * c, int batch_id, int n_tokens, int n_chunks, const uint8_t * descs, size_t desc_stride,
                      const KvOp * ops, int n_ops) {
    if (desc_stride != c->desc_stride) { pib_set_error("stage: descriptor stride %zu != %zu (contexts differ in max_seq)", desc_stride, c->desc_stride); return PIB_ERR_ARG; }
    if (n_tokens == 0) {                                  // KV ops only (pib_kv_sync_replicas on the head)
        std::vector<KvOp> & q = devq(c);
        q.insert(q.end(), ops, ops + n_ops);
        return pib_apply_devops(c);
    }
    pib_run * run = take_run(c, n_tokens);
    if (!run) return PIB_ERR_NOMEM;
    std::vector<int> slots(n_chunks);
    for (int ch = 0; ch < n_chunks; ch++) {
        const int s = take_desc_slot(c, run, n_chunks);
        slots[ch]   = s;
        uint8_t * d = c->desc_host + (size_t) s * c->desc_stride;
        memcpy(d, descs + (size_t) ch * desc_stride, desc_stride);
        cudaMemcpyAsync(c->desc_dev + (size_t) dev_slot(c, s) * c->desc_stride, d, c->desc_stride, cudaMemcpyHostToDevice, c->stream);
        c->h2d_bytes += c->desc_stride;
    }
    run->has_walk = n_chunks == 1 && reinterpret_cast<const int32_t *>(descs + PIB_DESC_HDR_OFF)[PIB_DH_TREE_NP] > 0;
    std::vector<KvOp> & dq = devq(c);
    for (int i = 0; i < n_ops; i++) {
        KvOp op = ops[i];
        if (op.op == KVOP_ALLOC) {
            if (op.p1 < 0 || op.p1 >= n_chunks) { pib_set_error("stage: bad chunk index in op-log"); return PIB_ERR_ARG; }
            op.p1 = slots[op.p1];
        }
        dq.push_back(op);
    }
    return enqueue_decode(c, run, n_tokens, batch_id, slots) >= 0 ? PIB_OK : PIB_ERR_CUDA;
}

// Clear a run slot's cancel flag in stream order after the run's last kernel -- and after any cancel write still travelling on the
// control stream: without the event a late cancel copy could land after the clear and poison the slot's next run.
static void clear_cancel_flag(pib_context * c, int slot) {
    if (c->cancel_ev_pending) { cudaStreamWaitEvent(c->stream, c->cancel_ev, 0); c->cancel_ev_pending = false; }
    cudaMemsetAsync(c->cancel_table + 1 + slot, 0, sizeof(int), c->stream);
}

void pib_stage_finish(pib_context * c, int batch_id) {
    for (auto it = c->inflight.begin(); it != c->inflight.end(); ++it) {
        pib_run & r = c->runs[*it];
        if (r.in_use && r.batch_id == batch_id) {
            clear_cancel_flag(c, r.id);
            r.in_use = false;
            c->last_finished = r.id;     // its logits stay reserved until the next finish, as on the head (llama_get_logits_ith)
            c->inflight.erase(it);
            return;
        }
    }
}

bool pib_run_ready(pib_context * c, int run) {
    if (run < 0 || run >= (int) c->runs.size() || !c->runs[run].in_use) return false;
    return cudaEventQuery(c->runs[run].done) == cudaSuccess;
}

int pib_run_finish(pib_context * c, int run) {
    if (run < 0 || run >= (int) c->runs.size() || !c->runs[run].in_use) { pib_set_error("finish: bad run handle"); return -1; }
    pib_run & r = c->runs[run];
    if (cudaEventSynchronize(r.done) != cudaSuccess) { pib_set_error("finish: %s", cudaGetErrorString(cudaGetLastError())); return -1; }
    // the run slot may be reused from here on: drop its cancel flag (ordered after every kernel of the run)
    clear_cancel_flag(c, r.id);
    if (c->stage.size > 1 && c->stage.first()) pib_pipe_ship_ints(2, 0, &r.batch_id, 1);
    for (auto it = c->inflight.begin(); it != c->inflight.end(); ++it) if (*it == run) { c->inflight.erase(it); break; }
    r.in_use         = false;
    c->last_finished = run;     // its buffers stay reserved until the next finish (llama_get_logits_ith contract)
    c->logits_row_valid.assign((size_t) r.n_tokens, 0);
    return r.failed ? -1 : 0;
}

const pib_row_result * pib_last_results(pib_context * c) {
    if (c->last_finished < 0) return nullptr;
    pib_run & r = c->runs[c->last_finished];
    if (!r.rows_on_host) {
        // a run whose acceptance was walked on the device: its rows were not copied; a caller that asks for them gets them now
        if (c->stage.size > 1) return nullptr;                 // they live on the last pipeline stage
        cudaMemcpyAsync(r.res_host, r.res_dev, sizeof(pib_row_result) * (size_t) r.n_tokens, cudaMemcpyDeviceToHost, c->stream);
        cudaStreamSynchronize(c->stream);
        c->d2h_bytes += sizeof(pib_row_result) * (size_t) r.n_tokens;
        r.rows_on_host = true;
    }
    return r.res_host;
}

const pib_accept_result * pib_last_accept(pib_context * c) {
    if (c->last_finished < 0 || !c->runs[c->last_finished].has_walk) return nullptr;
    return c->runs[c->last_finished].acc_host;
}

const float * pib_logits_row(pib_context * c, int i) {
    if (c->last_finished < 0) return nullptr;
    pib_run & r = c->runs[c->last_finished];
    const bool remote = !r.logits && c->stage.size > 1 && c->stage.first();       // the rows live on the last pipeline stage
    if (i < 0 || i >= r.n_tokens || (!r.logits && !remote)) return nullptr;
    const size_t V = (size_t) c->model->n_vocab;
    if (c->logits_host.size() < (size_t) r.n_tokens * V) c->logits_host.resize((size_t) r.n_tokens * V);
    if (!c->logits_row_valid[i]) {
        if (remote) {
            if (!c->logits_fetch_dev && cudaMalloc(&c->logits_fetch_dev, V * 4) != cudaSuccess) { cudaGetLastError(); return nullptr; }
            if (pib_pipe_fetch_logits(c, i, c->logits_fetch_dev, c->logits_host.data() + (size_t) i * V) != PIB_OK) return nullptr;
        } else {
            cudaMemcpyAsync(c->logits_host.data() + (size_t) i * V, r.logits + (size_t) i * V, V * 4, cudaMemcpyDeviceToHost, c->stream);
            cudaStreamSynchronize(c->stream);
            c->d2h_bytes += V * 4;
        }
        c->logits_row_valid[i] = 1;
    }
    return c->logits_host.data() + (size_t) i * V;
}

void pib_cancel(pib_context * c, const int32_t * batch_ids, int n) {
    // one cancel word per run slot (fixed addresses, so that captured graphs of a slot stay valid); the word is cleared when
    // the run is finished, in stream order after its last kernel
    bool wrote = false;
    for (int i = 0; i < n; i++)
        for (int id : c->inflight) {
            pib_run & r = c->runs[id];
            if (r.in_use && r.batch_id == batch_ids[i]) {
                cudaMemcpyAsync(c->cancel_table + 1 + r.id, c->cancel_one, sizeof(int), cudaMemcpyDeviceToDevice, c->ctrl);
                wrote = true;
            }
        }
    if (wrote) { cudaEventRecord(c->cancel_ev, c->ctrl); c->cancel_ev_pending = true; }
    if (c->stage.size > 1 && c->stage.first()) pib_pipe_ship_cancel(batch_ids, n);
}

void pib_set_logit_bias(pib_context * c, const int32_t * ids, const float * vals, int n) {
    if (c->stage.size > 1 && c->stage.first()) pib_pipe_ship_bias(ids, vals, n);
    cudaStreamSynchronize(c->stream);
    for (auto & g : c->graphs) if (g.second.exec) cudaGraphExecDestroy(g.second.exec);       // they hold the old bias pointers
    c->graphs.clear();
    if (c->bias_ids_dev) cudaFree(c->bias_ids_dev);
    if (c->bias_val_dev) cudaFree(c->bias_val_dev);
    c->bias_ids_dev = nullptr; c->bias_val_dev = nullptr; c->n_bias = 0;
    if (n <= 0) return;
    if (cudaMalloc(&c->bias_ids_dev, sizeof(int32_t) * n) != cudaSuccess || cudaMalloc(&c->bias_val_dev, sizeof(float) * n) != cudaSuccess ||
        cudaMemcpy(c->bias_ids_dev, ids, sizeof(int32_t) * n, cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy(c->bias_val_dev, vals, sizeof(float) * n, cudaMemcpyHostToDevice) != cudaSuccess) {
        // greedy output with the bias silently dropped would be wrong (e.g. --ignore-eos): fail loudly, as the reference's GGML_ASSERTs do
        fprintf(stderr, "pib_set_logit_bias: %s\n", cudaGetErrorString(cudaGetLastError()));
        abort();
    }
    c->n_bias = n;
}
