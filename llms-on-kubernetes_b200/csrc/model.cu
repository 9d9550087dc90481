// model.cu -- GGUF -> HBM upload and the decode step (SURVEY 8a rows a8-a16 end to end).
//
// Replaces, inside the process the chart launches (`llama-server --model P`,
// ramalama-models/helm-chart/templates/model-deployments.yaml:26-35), llama.cpp's model load and
// llama_decode for one token: embedding row -> L x {RMSNorm+QKV+RoPE+KV-write, attention,
// O-proj+residual, RMSNorm+gate/up+SwiGLU, down-proj+residual} -> RMSNorm+lm_head -> argmax.
//
// HBM layout: the GGUF data section is uploaded as ONE allocation, byte for byte, at the same
// relative offsets (so every tensor keeps its 32-byte file alignment and nothing is repacked); the
// KV cache is one paged fp16 pool [layer][kv head][pool row][head_dim] (model.h); activations are a
// handful of fp32 vectors.  Paths through this file:
//   one sequence, one token : the persistent step kernel (step.cu; one launch per token, replayed as a CUDA graph with
//                             token id, position and step counter on the device), or one launch per stage (B2_NO_STEP,
//                             mixed activation formats, tensor-parallel shards with an NCCL all-reduce per half-layer)
//   batches of columns      : b2_ctx_decode_batch -- int8 mma.sync kernel (gemm.cu) below 24 columns, the tensor-core
//                             pass (tc_gemm.cu) above, per-column positions / sequence slots, attn_batch.cu
//   prompts                 : b2_ctx_prefill(_multi) -- the tensor-core pass over whole chunks with causal attention
//                             (prefill_attn.cu); b2_ctx_decode takes it for prompts of 24+ tokens
// plus the resident fp16 copy of the model the tensor-core pass reads, the KV page table, and tensor-parallel shards.
#include "model.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>

#include <algorithm>

#include "gemm.cuh"
#include "gemv.cuh"
#include "step.cuh"
#include "ops.h"
#include "tc_gemm.cuh"
#include "tcq_gemm.cuh"

using namespace b2;

namespace {

bool env_flag(const char* name) {
    const char* v = getenv(name);
    return v && *v && strcmp(v, "0") != 0;
}

int find_w(b2_model* m, const std::string& name, b2_model::W* w, bool required) {
    const int64_t i = m->g->find_tensor(name);
    if (i < 0) {
        if (!required) return 1;
        set_error("model: tensor '%s' missing from '%s'", name.c_str(), m->g->path.c_str());
        return B2_EFORMAT;
    }
    const auto& t = m->g->tensors[i];
    w->p = m->d_data + t.offset;
    w->type = (int)t.type;
    w->cols = (int64_t)t.dims[0];
    w->rows = t.n_dims > 1 ? (int64_t)t.dims[1] : 1;
    return 0;
}

int check_matrix(const b2_model::W& w, const char* what, int64_t rows, int64_t cols) {
    if (w.rows != rows || w.cols != cols) {
        set_error("model: %s has shape [%lld x %lld], expected [%lld x %lld]", what, (long long)w.rows, (long long)w.cols,
                  (long long)rows, (long long)cols);
        return B2_EFORMAT;
    }
    if (gemv_act_for_type(w.type) < 0) {
        const TypeTraits* t = type_traits(w.type);
        set_error("model: %s is stored as ggml type %d (%s); this build computes Q4_K, Q5_K, Q6_K and Q8_0 matrices", what, w.type,
                  t ? t->name : "unknown");
        return B2_EUNSUP;
    }
    return 0;
}

int check_vec(const b2_model::W& w, const char* what, int64_t n) {
    if (w.type != B2_TYPE_F32 || w.cols * w.rows != n) {
        set_error("model: %s must be F32[%lld]", what, (long long)n);
        return B2_EFORMAT;
    }
    return 0;
}

// Tensor-parallel shard `rank` of `size`: attn_q/k/v by heads and ffn_gate/up by ff rows are row ranges of
// the matrices already in HBM (pointer + row count); attn_output and ffn_down are split along K on block boundaries --
// every row keeps blocks [rank * K/size, (rank + 1) * K/size), copied into a compact matrix.  The blocks themselves are
// untouched.  Norms, the embedding table and the lm_head stay whole on every rank; info.n_head / n_kv / d_ff become the
// shard's.  (The whole file image stays resident: 42 GB of the 180 on the 70B model; trimming it is bookkeeping.)
int shard_model(b2_model* m, int rank, int size) {
    b2_model_info& I = m->info;
    const int hd = I.head_dim;
    if (I.n_head % size || I.n_kv % size || I.d_ff % size) {
        set_error("model: %d heads / %d kv heads / d_ff %d do not divide over %d tensor-parallel ranks", I.n_head, I.n_kv, I.d_ff, size);
        return B2_EUNSUP;
    }
    const int64_t h_l = I.n_head / size, kv_l = I.n_kv / size, ff_l = I.d_ff / size;
    auto rows_of = [&](b2_model::W& w, int64_t first, int64_t n) {
        w.p += first * row_bytes(w.type, w.cols);
        w.rows = n;
    };
    auto k_slice = [&](b2_model::W& w, int64_t first, int64_t n, const char* what) -> int {
        const TypeTraits* t = type_traits(w.type);
        if (first % t->block_elems || n % t->block_elems || (gemv_act_for_type(w.type) == ACT_Q8K && n % 256)) {
            set_error("model: the K slice [%lld, +%lld) of %s does not fall on block boundaries", (long long)first, (long long)n, what);
            return B2_EUNSUP;
        }
        const size_t full = (size_t)row_bytes(w.type, w.cols), part = (size_t)row_bytes(w.type, n);
        const size_t off = (size_t)(first / t->block_elems) * t->block_bytes;
        uint8_t* dst = nullptr;
        B2_CUDA_CHECK(cudaMalloc((void**)&dst, part * w.rows + 256));
        m->tp_allocs.push_back(dst);
        B2_CUDA_CHECK(cudaMemcpy2D(dst, part, w.p + off, full, part, (size_t)w.rows, cudaMemcpyDeviceToDevice));
        B2_CUDA_CHECK(cudaMemset(dst + part * w.rows, 0, 256));
        w.p = dst;
        w.cols = n;
        return 0;
    };
    uint64_t stream = m->output.rows * (uint64_t)row_bytes(m->output.type, m->output.cols) + (uint64_t)I.d_model * 4;
    int rc;
    for (auto& L : m->layers) {
        rows_of(L.wq, rank * h_l * hd, h_l * hd);
        rows_of(L.wk, rank * kv_l * hd, kv_l * hd);
        rows_of(L.wv, rank * kv_l * hd, kv_l * hd);
        if (L.bq) L.bq += rank * h_l * hd;
        if (L.bk) L.bk += rank * kv_l * hd;
        if (L.bv) L.bv += rank * kv_l * hd;
        rows_of(L.gate, rank * ff_l, ff_l);
        rows_of(L.up, rank * ff_l, ff_l);
        if ((rc = k_slice(L.wo, rank * h_l * hd, h_l * hd, "attn_output.weight"))) return rc;
        if ((rc = k_slice(L.down, rank * ff_l, ff_l, "ffn_down.weight"))) return rc;
        const b2_model::W* ws[] = {&L.wq, &L.wk, &L.wv, &L.wo, &L.gate, &L.up, &L.down};
        for (const b2_model::W* w : ws) stream += (uint64_t)w->rows * row_bytes(w->type, w->cols);
        stream += 2ull * I.d_model * 4;
    }
    I.n_head = (int32_t)h_l;
    I.n_kv = (int32_t)kv_l;
    I.d_ff = (int32_t)ff_l;
    I.stream_bytes = stream;
    m->tp_rank = rank;
    m->tp_size = size;
    return 0;
}

}  // namespace

extern "C" int b2_model_load(const char* path, int device, int tp_rank, int tp_size, b2_model** out) {
    if (!out || !path) {
        set_error("b2_model_load: NULL argument");
        return B2_EINVAL;
    }
    *out = nullptr;
    if (tp_size < 1 || tp_size > 64 || tp_rank < 0 || tp_rank >= tp_size) {
        set_error("b2_model_load: tensor-parallel rank %d of %d is not valid", tp_rank, tp_size);
        return B2_EINVAL;
    }
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) {
        cudaGetLastError();
        set_error("b2_model_load: no CUDA device visible; this library has no CPU fallback");
        return B2_ECUDA;
    }
    if (device < 0 || device >= n_dev) {
        set_error("b2_model_load: device %d out of range (0..%d)", device, n_dev - 1);
        return B2_EINVAL;
    }
    B2_CUDA_CHECK(cudaSetDevice(device));
    Gguf* gp = nullptr;
    int rc = Gguf::open(path, &gp);
    if (rc) return rc;
    std::unique_ptr<b2_model> m(new b2_model());
    m->g.reset(gp);
    m->gguf_handle.g = gp;
    m->gguf_handle.owned = false;
    m->device = device;
    Gguf& g = *gp;

    std::string arch;
    if (!g.get_str("general.architecture", &arch)) {
        set_error("model: general.architecture missing");
        return B2_EFORMAT;
    }
    b2_model_info& I = m->info;
    if (arch == "llama") I.arch = 0;
    else if (arch == "qwen2") I.arch = 1;
    else {
        set_error("model: architecture '%s' is not supported (llama, qwen2)", arch.c_str());
        return B2_EUNSUP;
    }
    uint32_t u;
    auto need_u32 = [&](const std::string& k, int32_t* dst) -> int {
        if (!g.get_u32(arch + "." + k, &u)) {
            set_error("model: key %s.%s missing", arch.c_str(), k.c_str());
            return B2_EFORMAT;
        }
        *dst = (int32_t)u;
        return 0;
    };
    if ((rc = need_u32("block_count", &I.n_layer))) return rc;
    if ((rc = need_u32("embedding_length", &I.d_model))) return rc;
    if ((rc = need_u32("feed_forward_length", &I.d_ff))) return rc;
    if ((rc = need_u32("attention.head_count", &I.n_head))) return rc;
    I.n_kv = g.get_u32(arch + ".attention.head_count_kv", &u) ? (int32_t)u : I.n_head;
    if (I.n_head <= 0 || I.n_kv <= 0 || I.n_layer <= 0 || I.d_model <= 0 || I.d_ff <= 0) {
        set_error("model: non-positive hyper-parameter");
        return B2_EFORMAT;
    }
    I.head_dim = g.get_u32(arch + ".attention.key_length", &u) ? (int32_t)u : I.d_model / I.n_head;
    I.rope_dim = g.get_u32(arch + ".rope.dimension_count", &u) ? (int32_t)u : I.head_dim;
    I.n_ctx_train = g.get_u32(arch + ".context_length", &u) ? (int32_t)u : 2048;
    if (!g.get_f32(arch + ".attention.layer_norm_rms_epsilon", &I.eps)) I.eps = 1e-5f;
    if (!g.get_f32(arch + ".rope.freq_base", &I.rope_base)) I.rope_base = 10000.0f;
    I.bos_id = g.get_u32("tokenizer.ggml.bos_token_id", &u) ? (int32_t)u : -1;
    I.eos_id = g.get_u32("tokenizer.ggml.eos_token_id", &u) ? (int32_t)u : -1;
    I.eot_id = g.get_u32("tokenizer.ggml.eot_token_id", &u) ? (int32_t)u : -1;
    if (I.n_head % I.n_kv || I.rope_dim > I.head_dim || (I.rope_dim & 1)) {
        set_error("model: inconsistent head configuration");
        return B2_EFORMAT;
    }

    // ---- upload the data section as-is: pinned double-buffered staging from the mmap ----
    const size_t guard = 256;
    B2_CUDA_CHECK(cudaMalloc((void**)&m->d_alloc, g.data_size + 2 * guard));
    m->d_data = m->d_alloc + guard;
    B2_CUDA_CHECK(cudaMemset(m->d_alloc, 0, guard));
    B2_CUDA_CHECK(cudaMemset(m->d_data + g.data_size, 0, guard));
    {
        const size_t CH = 64u << 20;
        uint8_t* pin[2] = {nullptr, nullptr};
        cudaEvent_t ev[2];
        cudaStream_t st;
        B2_CUDA_CHECK(cudaStreamCreate(&st));
        for (int i = 0; i < 2; i++) {
            B2_CUDA_CHECK(cudaMallocHost((void**)&pin[i], CH));
            B2_CUDA_CHECK(cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming));
        }
        const uint8_t* src = g.map + g.data_offset;
        int b = 0;
        for (size_t off = 0; off < g.data_size; off += CH, b ^= 1) {
            const size_t n = std::min(CH, (size_t)g.data_size - off);
            B2_CUDA_CHECK(cudaEventSynchronize(ev[b]));
            memcpy(pin[b], src + off, n);
            B2_CUDA_CHECK(cudaMemcpyAsync(m->d_data + off, pin[b], n, cudaMemcpyHostToDevice, st));
            B2_CUDA_CHECK(cudaEventRecord(ev[b], st));
        }
        B2_CUDA_CHECK(cudaStreamSynchronize(st));
        for (int i = 0; i < 2; i++) {
            cudaFreeHost(pin[i]);
            cudaEventDestroy(ev[i]);
        }
        cudaStreamDestroy(st);
    }
    I.weight_bytes = g.data_size;

    // ---- tensor table ----
    if ((rc = find_w(m.get(), "token_embd.weight", &m->tok_embd, true)) < 0) return rc;
    if ((rc = find_w(m.get(), "output_norm.weight", &m->output_norm, true)) < 0) return rc;
    rc = find_w(m.get(), "output.weight", &m->output, false);
    if (rc < 0) return rc;
    I.tied = rc == 1;
    if (I.tied) m->output = m->tok_embd;
    I.vocab = (int32_t)m->tok_embd.rows;
    if (m->tok_embd.cols != I.d_model || !type_traits(m->tok_embd.type)) {
        set_error("model: token_embd.weight has an unsupported shape or type");
        return B2_EUNSUP;
    }
    if ((rc = check_matrix(m->output, "output.weight", I.vocab, I.d_model))) return rc;
    if ((rc = check_vec(m->output_norm, "output_norm.weight", I.d_model))) return rc;
    {
        b2_model::W rf;
        rc = find_w(m.get(), "rope_freqs.weight", &rf, false);
        if (rc < 0) return rc;
        if (rc == 0) {
            if ((rc = check_vec(rf, "rope_freqs.weight", I.rope_dim / 2))) return rc;
            m->rope_freqs = (const float*)rf.p;
            const int64_t ti = g.find_tensor("rope_freqs.weight");
            const float* hp = (const float*)g.tensor_data(ti);
            m->rope_freqs_host.assign(hp, hp + I.rope_dim / 2);
        }
    }
    const int64_t qd = (int64_t)I.n_head * I.head_dim, kvd = (int64_t)I.n_kv * I.head_dim;
    uint64_t stream = m->output.rows * (uint64_t)row_bytes(m->output.type, m->output.cols) + (uint64_t)I.d_model * 4;
    m->layers.resize(I.n_layer);
    for (int il = 0; il < I.n_layer; il++) {
        auto& L = m->layers[il];
        const std::string p = "blk." + std::to_string(il) + ".";
        struct { const char* n; b2_model::W* w; int64_t r, c; } mats[] = {
            {"attn_q.weight", &L.wq, qd, I.d_model},   {"attn_k.weight", &L.wk, kvd, I.d_model},
            {"attn_v.weight", &L.wv, kvd, I.d_model},  {"attn_output.weight", &L.wo, I.d_model, qd},
            {"ffn_gate.weight", &L.gate, I.d_ff, I.d_model}, {"ffn_up.weight", &L.up, I.d_ff, I.d_model},
            {"ffn_down.weight", &L.down, I.d_model, I.d_ff}};
        for (auto& e : mats) {
            if ((rc = find_w(m.get(), p + e.n, e.w, true)) < 0) return rc;
            if ((rc = check_matrix(*e.w, (p + e.n).c_str(), e.r, e.c))) return rc;
            stream += (uint64_t)e.w->rows * row_bytes(e.w->type, e.w->cols);
        }
        if ((rc = find_w(m.get(), p + "attn_norm.weight", &L.attn_norm, true)) < 0) return rc;
        if ((rc = find_w(m.get(), p + "ffn_norm.weight", &L.ffn_norm, true)) < 0) return rc;
        if ((rc = check_vec(L.attn_norm, "attn_norm.weight", I.d_model))) return rc;
        if ((rc = check_vec(L.ffn_norm, "ffn_norm.weight", I.d_model))) return rc;
        stream += 2ull * I.d_model * 4;
        struct { const char* n; const float** b; int64_t len; } biases[] = {
            {"attn_q.bias", &L.bq, qd}, {"attn_k.bias", &L.bk, kvd}, {"attn_v.bias", &L.bv, kvd}};
        for (auto& e : biases) {
            b2_model::W bw;
            rc = find_w(m.get(), p + e.n, &bw, false);
            if (rc < 0) return rc;
            if (rc == 0) {
                if ((rc = check_vec(bw, e.n, e.len))) return rc;
                *e.b = (const float*)bw.p;
                stream += (uint64_t)e.len * 4;
            }
        }
    }
    I.stream_bytes = stream;
    if (tp_size > 1 && (rc = shard_model(m.get(), tp_rank, tp_size))) return rc;
    *out = m.release();
    return 0;
}

extern "C" void b2_model_free(b2_model* m) {
    if (!m) return;
    if (m->d_alloc) cudaFree(m->d_alloc);
    if (m->f16_cache) cudaFree(m->f16_cache);
    for (void* p : m->tp_allocs) cudaFree(p);
    delete m;
}

extern "C" int b2_model_get_info(const b2_model* m, b2_model_info* out) {
    if (!m || !out) return B2_EINVAL;
    *out = m->info;
    return 0;
}
extern "C" const b2_gguf* b2_model_gguf(const b2_model* m) { return m ? &m->gguf_handle : nullptr; }

extern "C" int b2_model_tensor(const b2_model* m, const char* name, const void** dev_ptr, b2_tensor_info* info) {
    if (!m || !name) return B2_EINVAL;
    const int64_t i = m->g->find_tensor(name);
    if (i < 0) {
        set_error("b2_model_tensor: no tensor '%s'", name);
        return B2_ENOTFOUND;
    }
    const auto& t = m->g->tensors[i];
    if (dev_ptr) *dev_ptr = m->d_data + t.offset;
    if (info) {
        info->name = t.name.c_str();
        info->n_dims = t.n_dims;
        for (int d = 0; d < 4; d++) info->dims[d] = t.dims[d];
        info->type = t.type;
        info->offset = t.offset;
        info->n_bytes = t.n_bytes;
    }
    return 0;
}

// -------------------------------------------------------------------------------------------------
// decode context
// -------------------------------------------------------------------------------------------------
namespace {

template <typename T>
int dmalloc(T** p, size_t n) {
    B2_CUDA_CHECK(cudaMalloc((void**)p, n * sizeof(T)));
    return 0;
}

void seg_matrix(GemvSeg& s, const b2_model::W& w, int pairing, int epi) {
    s = GemvSeg{};
    s.wa = w.p;
    s.type = w.type;
    s.pairing = pairing;
    s.epi = epi;
    s.n_rows = (int)w.rows;
    s.n_pairs = pairing == PAIR_SPLIT ? (int)w.rows : (int)((w.rows + 1) / 2);
}

// enqueue one decode step (token and position are read from device memory)
// With `record` set nothing is launched: the stages are appended to *record as the program of the persistent
// per-token kernel (step.cu), and *record_act receives the activation format they share (-2 if they do not).
int enqueue_step(b2_ctx* c, cudaStream_t st, bool pdl, bool matvec_only = false, std::vector<StepOp>* record = nullptr,
                 int* record_act = nullptr) {
    b2_model* m = c->m;
    const b2_model_info& I = m->info;
    int rc;
    int launches = 0;
    if (record_act) *record_act = -1;
    const int64_t kv_layer = (int64_t)I.n_kv * c->kv_tokens * I.head_dim;   // slot 0 = rows [0, n_ctx) of every kv head
    const int rope_pairing = I.arch == 1 ? PAIR_NEOX : PAIR_ADJ;
    auto common = [&](GemvArgs& a, const float* x, const float* norm_w, int K) {
        a = GemvArgs{};
        a.x = x;
        a.norm_w = norm_w;
        a.prologue = norm_w ? PRO_NORM : PRO_PLAIN;
        a.eps = I.eps;
        a.K = K;
        a.pos_ptr = c->d_pos;
        a.rope_tab = c->rope_tab;
        a.rope_half = I.rope_dim / 2;
    };
    auto run = [&](GemvArgs& a, int act) -> int {
        if ((rc = gemv_finalize(a, act))) return rc;
        if (record) {
            StepOp op{};
            op.kind = OP_GEMV;
            op.g = a;
            record->push_back(op);
            if (*record_act == -1) *record_act = act;
            else if (*record_act != act) *record_act = -2;
            return 0;
        }
        if ((rc = launch_gemv(a, act, st, pdl))) return rc;
        launches++;
        return 0;
    };
    // the first kernel of a step depends on a memcpy / the previous graph launch: ordinary dependency
    if (record) {
        StepOp op{};
        op.kind = OP_EMBED;
        op.e_type = m->tok_embd.type;
        op.e_table = m->tok_embd.p;
        op.e_row_bytes = row_bytes(m->tok_embd.type, I.d_model);
        op.e_d_model = I.d_model;
        op.e_vocab = I.vocab;
        op.e_token = c->d_tok;
        op.e_x = c->x;
        record->push_back(op);
    } else if (!matvec_only) {
        if ((rc = launch_embed(m->tok_embd.type, m->tok_embd.p, row_bytes(m->tok_embd.type, I.d_model), I.d_model, I.vocab, c->d_tok, c->x,
                               st, false)))
            return rc;
        launches++;
    }
    for (int il = 0; il < I.n_layer; il++) {
        const auto& L = m->layers[il];
        __half* kc = c->k_cache + il * kv_layer;
        __half* vc = c->v_cache + il * kv_layer;
        // --- attention block: RMSNorm -> Q,K,V (+bias) -> RoPE -> cache ---
        GemvSeg sq, sk, sv;
        seg_matrix(sq, L.wq, rope_pairing, EPI_ROPE_Q);
        sq.out = c->q;
        sq.bias = L.bq;
        sq.head_dim = I.head_dim;
        seg_matrix(sk, L.wk, rope_pairing, EPI_ROPE_K);
        sk.cache = kc;
        sk.bias = L.bk;
        sk.head_dim = I.head_dim;
        sk.n_ctx = c->kv_tokens;
        seg_matrix(sv, L.wv, PAIR_ADJ, EPI_V);
        sv.cache = vc;
        sv.bias = L.bv;
        sv.head_dim = I.head_dim;
        sv.n_ctx = c->kv_tokens;
        const int aq = gemv_act_for_type(L.wq.type), ak = gemv_act_for_type(L.wk.type), av = gemv_act_for_type(L.wv.type);
        GemvArgs a;
        if (aq == ak && ak == av) {
            common(a, c->x, (const float*)L.attn_norm.p, I.d_model);
            a.seg[0] = sq; a.seg[1] = sk; a.seg[2] = sv;
            a.n_seg = 3;
            if ((rc = run(a, aq))) return rc;
        } else {
            const GemvSeg* segs[3] = {&sq, &sk, &sv};
            const int acts[3] = {aq, ak, av};
            for (int i = 0; i < 3; i++) {
                common(a, c->x, (const float*)L.attn_norm.p, I.d_model);
                a.seg[0] = *segs[i];
                a.n_seg = 1;
                if ((rc = run(a, acts[i]))) return rc;
            }
        }
        if (record) {
            StepOp op{};
            op.kind = OP_ATTN;
            op.a_q = c->q;
            op.a_k = kc;
            op.a_v = vc;
            op.a_out = c->attn_out;
            op.a_n_head = I.n_head;
            op.a_n_kv = I.n_kv;
            op.a_hd = I.head_dim;
            op.a_n_ctx = c->kv_tokens;   // row stride of a kv head
            op.a_pos = c->d_pos;   // a_max_chunks / a_partials: build_step_program(), once the grid is known
            op.a_counters = c->step_counters + 8;   // [<= n_head] after the grid-barrier and argmax counters
            record->push_back(op);
        } else if (!matvec_only) {
            if ((rc = launch_attn_decode(c->q, kc, vc, c->attn_out, I.n_head, I.n_kv, I.head_dim, c->kv_tokens, c->d_pos, 0, c->max_splits,
                                         c->attn_scratch, st, pdl)))
                return rc;
            launches++;
        }
        // --- attention output projection + residual ---
        // (tensor parallel: this rank's K slice gives a partial vector -> summed over the ranks, then added to x)
        const bool tp = m->tp_size > 1;
        common(a, c->attn_out, nullptr, I.n_head * I.head_dim);
        seg_matrix(a.seg[0], L.wo, PAIR_ADJ, tp ? EPI_RESID_TP : EPI_RESID);
        a.seg[0].out = c->x;
        a.seg[0].res_scale = 1.0f / (float)m->tp_size;   // exact for 2 / 4 / 8 ranks
        a.n_seg = 1;
        if ((rc = run(a, gemv_act_for_type(L.wo.type)))) return rc;
        if (tp && !record && !matvec_only) {
            if ((rc = tp_allreduce_inplace(c->tp_comm, c->x, I.d_model, st))) return rc;
            launches += 1;
        }
        // --- FFN: RMSNorm -> gate, up -> SwiGLU ---
        const int ag = gemv_act_for_type(L.gate.type), au = gemv_act_for_type(L.up.type);
        if (ag == au && L.gate.type == L.up.type) {
            common(a, c->x, (const float*)L.ffn_norm.p, I.d_model);
            seg_matrix(a.seg[0], L.gate, PAIR_SPLIT, EPI_SWIGLU);
            a.seg[0].wb = L.up.p;
            a.seg[0].out = c->h;
            a.n_seg = 1;
            if ((rc = run(a, ag))) return rc;
        } else {
            set_error("model: ffn_gate and ffn_up of layer %d have different types (%d, %d): unsupported", il, L.gate.type, L.up.type);
            return B2_EUNSUP;
        }
        // --- FFN down projection + residual ---
        common(a, c->h, nullptr, I.d_ff);
        seg_matrix(a.seg[0], L.down, PAIR_ADJ, tp ? EPI_RESID_TP : EPI_RESID);
        a.seg[0].out = c->x;
        a.seg[0].res_scale = 1.0f / (float)m->tp_size;
        a.n_seg = 1;
        if ((rc = run(a, gemv_act_for_type(L.down.type)))) return rc;
        if (tp && !record && !matvec_only) {
            if ((rc = tp_allreduce_inplace(c->tp_comm, c->x, I.d_model, st))) return rc;
            launches += 1;
        }
    }
    // --- final norm + lm_head ---
    GemvArgs a;
    common(a, c->x, (const float*)m->output_norm.p, I.d_model);
    seg_matrix(a.seg[0], m->output, PAIR_ADJ, EPI_STORE);
    a.seg[0].out = c->logits;
    a.n_seg = 1;
    if ((rc = run(a, gemv_act_for_type(m->output.type)))) return rc;
    if (record) {
        StepOp op{};
        op.kind = OP_ARGMAX;
        op.s_logits = c->logits;
        op.s_n = I.vocab;
        op.s_token = c->d_tok;
        op.s_pos = c->d_pos;
        op.s_step = c->d_step;
        op.s_out = c->d_out;
        op.s_cap = c->out_cap;
        op.s_val = c->step_argmax_val;
        op.s_idx = c->step_argmax_idx;
        op.s_counter = c->step_counters + 1;
        record->push_back(op);
        return 0;
    }
    if (matvec_only) {
        c->matvec_launches = launches;
        return 0;
    }
    if ((rc = launch_argmax(c->logits, I.vocab, c->d_tok, c->argmax_scratch, c->d_pos, c->d_step, c->d_out, c->out_cap, st, pdl))) return rc;
    launches++;
    c->launches_per_step = launches;
    return 0;
}

bool slot_samples(const b2_ctx* c, int slot) {
    return c->n_samplers > 0 && slot >= 0 && slot < (int)c->samplers.size() && c->samplers[slot].temperature > 0.0f && c->samplers[slot].top_k != 1;
}

// Slot 0 samples: replace the greedy token the step just left in d_tok (and, inside the device-resident loop, in the
// output list) by a draw from the step's logits.  The draw index is the position the new token will take (d_pos after
// the step).  after_step: the step advanced d_step, so the list entry to overwrite is d_out[*d_step - 1].
int sample_single(b2_ctx* c, cudaStream_t st, bool after_step) {
    if (!slot_samples(c, 0)) return 0;
    return launch_sample_rows(c->logits, c->m->info.vocab, 0, 1, c->d_samplers, 0, nullptr, c->d_tok, true, st, c->d_pos,
                              after_step ? c->d_out : nullptr, after_step ? c->d_step : nullptr, c->out_cap);
}

// The columns of a batched pass whose sequence slots sample: one draw per logits row, on the device.  Row r of
// bt.logits belongs to column col0 + r of the pass (slot / position from the host arrays of the call).
int sample_columns(b2_ctx* c, const int32_t* slots, const int32_t* pos, int col0, int n_rows) {
    if (c->n_samplers == 0) return 0;
    bool any = false;
    for (int r = 0; r < n_rows; r++) any = any || slot_samples(c, slots ? slots[col0 + r] : 0);
    if (!any) return 0;
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));   // pinned staging reuse
    for (int r = 0; r < n_rows; r++) {
        const int s = slots ? slots[col0 + r] : 0;
        c->h_col_params[r] = slot_samples(c, s) ? c->samplers[s] : b2_sampling{0.0f, 1.0f, 0.0f, 1, 0};
        c->h_col_draw[r] = (uint64_t)(pos[col0 + r] + 1);
    }
    B2_CUDA_CHECK(cudaMemcpyAsync(c->d_col_params, c->h_col_params, (size_t)n_rows * sizeof(b2_sampling), cudaMemcpyHostToDevice, c->stream));
    B2_CUDA_CHECK(cudaMemcpyAsync(c->d_col_draw, c->h_col_draw, (size_t)n_rows * sizeof(uint64_t), cudaMemcpyHostToDevice, c->stream));
    return launch_sample_rows(c->bt.logits, c->m->info.vocab, c->m->info.vocab, n_rows, c->d_col_params, 1, c->d_col_draw, c->bt.out, true, c->stream);
}

// one decode step on the stream: the persistent kernel when a program was built, else one launch per stage
int enqueue_decode(b2_ctx* c, cudaStream_t st, bool pdl) {
    if (c->use_step && c->d_ops) {
        StepProgram p{};
        p.ops = (const StepOp*)c->d_ops;
        p.n_ops = c->n_ops;
        p.max_k = c->step_max_k;
        p.sync_counter = c->step_sync;
        p.trace = c->step_trace;
        p.l2_ahead = c->step_l2_ahead;
        p.fine_op = c->step_fine_op;
        p.tp = c->tp_local ? c->m->tp_size : 0;
        int rc = c->step_version == 2 ? launch_step2(p, c->step_act, st) : launch_step(p, c->step_act, st);
        if (rc == 0) c->launches_per_step = 1;
        if (rc == 0) rc = sample_single(c, st, true);
        return rc;
    }
    const int rc = enqueue_step(c, st, pdl);
    return rc ? rc : sample_single(c, st, true);
}

// record the stages of a step as the persistent kernel's program and upload it
int build_step_program(b2_ctx* c) {
    std::vector<StepOp> ops;
    int act = -1;
    int rc = enqueue_step(c, c->stream, false, false, &ops, &act);
    if (rc) return rc;
    if (act < 0) {
        // mat-vecs with different activation formats (K-quants mixed with Q8_0): keep one launch per stage
        c->use_step = false;
        return 0;
    }
    int max_k = 0;
    for (auto& op : ops)
        if (op.kind == OP_GEMV && op.g.K > max_k) max_k = op.g.K;
    c->step_version = 2;
    if (const char* e = getenv("B2_STEP_V")) c->step_version = atoi(e) == 1 ? 1 : 2;
    for (auto& op : ops)
        if (op.kind == OP_GEMV && !step2_supports(act, op.g.K)) c->step_version = 1;
    c->step_l2_ahead = 0;
    if (const char* e = getenv("B2_STEP_L2_AHEAD")) c->step_l2_ahead = atoi(e) < 0 ? 0 : atoi(e);
    const int grid = c->step_version == 2 ? step2_grid_size(act, max_k, c->tp_local != nullptr) : step_grid_size(act, max_k);
    if (grid < 0) return grid;
    if (grid > 1024) {
        set_error("step kernel: grid of %d CTAs exceeds the argmax scratch", grid);
        return B2_ERANGE;
    }
    // attention scratch depends on the grid (the chunk length is chosen so that the items cover it once)
    const auto& I = c->m->info;
    B2_CUDA_CHECK(cudaMalloc((void**)&c->step_attn_partials, step_attn_partials_bytes(I.n_head, I.head_dim, c->n_ctx, grid)));
    for (auto& op : ops)
        if (op.kind == OP_ATTN) {
            op.a_max_chunks = step_attn_max_chunks(c->n_ctx, grid);
            op.a_partials = c->step_attn_partials;
            op.a_min_chunk = STEP_ATTN_MIN_CHUNK;
            op.a_cluster_merge = env_flag("B2_ATTN_NO_CLUSTER") ? 0 : 1;
            if (const char* e = getenv("B2_ATTN_MIN_CHUNK")) {   // tuning knob: 32 .. 256 in steps of 32
                const int v = atoi(e);
                if (v >= 32 && v <= STEP_ATTN_MAX_CHUNK && v % 32 == 0) op.a_min_chunk = v;
            }
        }
    if (c->m->tp_size > 1) {
        // tensor-parallel shard: the exchange stages and the peer stores of the mat-vecs in front of them
        const int tp = c->m->tp_size, rank = c->m->tp_rank, d = I.d_model;
        if (c->step_version != 2 || !c->tp_local) {
            c->use_step = false;   // (the first design has no exchange stage)
            return 0;
        }
        // every row-split mat-vec (O, down) is exchange k: its epilogues store into all ranks' inboxes (parity k & 1), and the
        // next mat-vec stage -- the consumer of the new residual stream -- waits for the ranks' sequence numbers and reads the sum
        // B2_TP_GATHER=1: no reduce stage -- the consuming mat-vec's prologue adds up the ranks' rows itself (first version)
        const bool gather = env_flag("B2_TP_GATHER");
        // B2_TP_PUSH=1: the epilogues store their partial rows into every rank's inbox (system-scope fence per CTA); default:
        // the rows stay local and the reduce stage of every rank fetches its slice of them over NVLink
        const bool pull = !gather && !env_flag("B2_TP_PUSH");
        int k = 0;
        std::vector<StepOp> out;
        out.reserve(ops.size() + 2 * I.n_layer);
        for (size_t i = 0; i < ops.size(); i++) {
            StepOp& op = ops[i];
            if (op.kind == OP_ARGMAX) op.s_epoch = c->tp_epoch;
            if (op.kind != OP_GEMV || op.g.n_seg != 1 || op.g.seg[0].epi != EPI_RESID_TP) {
                out.push_back(op);
                continue;
            }
            size_t j = i + 1;
            while (j < ops.size() && ops[j].kind != OP_GEMV) j++;
            if (k >= 255 || j == ops.size() || ops[j].g.x != c->x || ops[j].g.K != d) {
                set_error("step program: row-split mat-vec %d is not followed by a stage that reads the residual stream", k);
                return B2_EINVAL;
            }
            GemvSeg& s = op.g.seg[0];
            if (pull) {
                // the partial row stays in this rank's buffer (slot `rank` of parity k & 1); the peers' reduce stages fetch it
                s.n_peer = 1;
                s.peer_local = 1;
                s.peer_out[0] = (float*)c->tp_local + ((size_t)(k & 1) * tp + rank) * d;
            } else {
                s.n_peer = tp;
                for (int p = 0; p < tp; p++) s.peer_out[p] = (float*)c->tp_peers[p] + ((size_t)(k & 1) * tp + rank) * d;
            }
            out.push_back(op);
            StepOp xr{};
            xr.kind = OP_XCHG;
            GemvArgs& g = gather ? ops[j].g : xr.g;
            g.x = c->x;
            g.K = d;
            g.x_parts = (const float*)c->tp_local + (size_t)(k & 1) * tp * d;
            g.x_nparts = tp;
            g.xg_k = k;
            g.xg_flags = (const unsigned int*)((const uint8_t*)c->tp_local + c->tp_flags_off);
            for (int p = 0; p < tp; p++) g.xg_peer_flags[p] = (unsigned int*)((uint8_t*)c->tp_peers[p] + c->tp_flags_off) + rank;
            g.xg_epoch = c->tp_epoch;
            if (pull)
                for (int p = 0; p < tp; p++) g.x_part_ptr[p] = (const float*)c->tp_peers[p] + ((size_t)(k & 1) * tp + p) * d;
            if (!gather) out.push_back(xr);
            k++;
        }
        ops.swap(out);
    }
    B2_CUDA_CHECK(cudaMalloc(&c->d_ops, ops.size() * sizeof(StepOp)));
    B2_CUDA_CHECK(cudaMemcpyAsync(c->d_ops, ops.data(), ops.size() * sizeof(StepOp), cudaMemcpyHostToDevice, c->stream));
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    c->n_ops = (int)ops.size();
    c->step_act = act;
    c->step_max_k = max_k;
    if (env_flag("B2_STEP_TRACE")) {
        const size_t words = ops.size() * STEP_TRACE_SLOTS + 8 * STEP_FINE_STEPS;
        B2_CUDA_CHECK(cudaMalloc((void**)&c->step_trace, words * sizeof(unsigned long long)));
        B2_CUDA_CHECK(cudaMemset(c->step_trace, 0, words * sizeof(unsigned long long)));
        if (const char* e = getenv("B2_STEP_FINE")) c->step_fine_op = atoi(e);   // -2: the compute warps skip the arithmetic (wrong results)
        for (auto& op : ops) {
            c->step_desc.push_back(op.kind);
            c->step_desc.push_back(op.kind == OP_GEMV ? op.g.K : 0);
            c->step_desc.push_back(op.kind == OP_GEMV ? op.g.n_tiles : 0);
        }
    }
    return 0;
}

// B2_STEP_TRACE=1: print where CTA 0 spent the last step, stage by stage (first two layers) and by stage class
void dump_step_trace(b2_ctx* c) {
    if (!c->step_trace) return;
    const int S = STEP_TRACE_SLOTS;
    std::vector<unsigned long long> t((size_t)c->n_ops * S);
    if (cudaMemcpy(t.data(), c->step_trace, t.size() * 8, cudaMemcpyDeviceToHost) != cudaSuccess) return;
    const char* kinds[] = {"embed", "gemv", "attn", "argmax", "xchg"};
    struct Acc { double wait = 0, pro = 0, tiles = 0, epi = 0, mark[7] = {0, 0, 0, 0, 0, 0, 0}; int n = 0; };
    std::map<std::string, Acc> by;
    unsigned long long prev_arrive = 0;
    fprintf(stderr, "b2 step trace (CTA 0, ns): stage kind K tiles | barrier-wait  prologue  tiles  epilogue+arrive | marks 4..7 after stage start\n");
    for (int i = 0; i < c->n_ops; i++) {
        const int kind = c->step_desc[3 * i], K = c->step_desc[3 * i + 1], tiles = c->step_desc[3 * i + 2];
        const unsigned long long* s = &t[(size_t)i * S];
        const double wait = prev_arrive ? (double)(s[0] - prev_arrive) : 0.0;
        const bool four = kind == OP_GEMV || kind == OP_ATTN;   // stages that stamp all four slots
        const double pro = four ? (double)(s[1] - s[0]) : 0.0;
        const double tl = four ? (double)(s[2] - s[1]) : 0.0;
        const double epi = (double)((long long)s[3] - (long long)(four ? s[2] : s[0]));
        char key[64];
        snprintf(key, sizeof key, "%s K=%d tiles=%d", kinds[kind], K, tiles);
        Acc& a = by[key];
        a.wait += wait; a.pro += pro; a.tiles += tl; a.epi += epi; a.n++;
        for (int k = 0; k < 7; k++) a.mark[k] += s[4 + k] ? (double)(s[4 + k] - s[0]) : 0.0;
        if (i < 12) fprintf(stderr, "  %3d %-28s | %8.0f %8.0f %8.0f %8.0f\n", i, key, wait, pro, tl, epi);
        prev_arrive = s[3];
    }
    fprintf(stderr, "b2 step trace, mean per stage class:\n");
    for (auto& kv : by) {
        const Acc& a = kv.second;
        fprintf(stderr, "  %-28s x%3d | %8.0f %8.0f %8.0f %8.0f | %6.0f %6.0f %6.0f %6.0f | ready %6.0f dot %6.0f free %6.0f\n", kv.first.c_str(), a.n, a.wait / a.n,
                a.pro / a.n, a.tiles / a.n, a.epi / a.n, a.mark[0] / a.n, a.mark[1] / a.n, a.mark[2] / a.n, a.mark[3] / a.n, a.mark[4] / a.n, a.mark[5] / a.n,
                a.mark[6] / a.n);
    }
    fprintf(stderr, "  whole step: %.0f ns\n", (double)(t[(size_t)(c->n_ops - 1) * S + 3] - t[0]));
    if (c->step_fine_op >= 0 && c->step_fine_op < c->n_ops) {
        std::vector<unsigned long long> f(8 * STEP_FINE_STEPS);
        if (cudaMemcpy(f.data(), c->step_trace + (size_t)c->n_ops * S, f.size() * 8, cudaMemcpyDeviceToHost) != cudaSuccess) return;
        {
            // arrival of every CTA on the traced stage's barrier and the moment it saw the barrier complete
            unsigned long long first = ~0ull, last = 0, pfirst = ~0ull, plast = 0;
            int n = 0;
            for (int b = 0; b < 1024 && f[1024 + b]; b++, n++) {
                first = std::min(first, f[1024 + b]); last = std::max(last, f[1024 + b]);
                if (f[2048 + b]) { pfirst = std::min(pfirst, f[2048 + b]); plast = std::max(plast, f[2048 + b]); }
            }
            if (n) {
                // deciles of both distributions (ns after the first arrival / after the last arrival)
                std::vector<long long> arr, det;
                for (int b = 0; b < n; b++) {
                    arr.push_back((long long)(f[1024 + b] - first));
                    if (f[2048 + b]) det.push_back((long long)f[2048 + b] - (long long)last);
                }
                std::sort(arr.begin(), arr.end());
                std::sort(det.begin(), det.end());
                fprintf(stderr, "b2 step trace, stage %d barrier, arrival deciles:", c->step_fine_op);
                for (int q = 0; q <= 10; q++) fprintf(stderr, " %lld", arr[(size_t)q * (arr.size() - 1) / 10]);
                fprintf(stderr, "\nb2 step trace, stage %d barrier, seen-complete deciles:", c->step_fine_op);
                for (int q = 0; q <= 10 && !det.empty(); q++) fprintf(stderr, " %lld", det[(size_t)q * (det.size() - 1) / 10]);
                fprintf(stderr, "\n");
            }
            if (n)
                fprintf(stderr, "b2 step trace, stage %d barrier: %d CTAs arrive over %lld ns (CTA 0 at +%lld); the barrier is seen complete %lld .. %lld ns after the last arrival\n",
                        c->step_fine_op, n, (long long)(last - first), (long long)(f[1024] - first), (long long)(pfirst - last), (long long)(plast - last));
        }
        const long long t0 = (long long)f[6];   // SM clock when CTA 0's compute warps passed the stage's barrier
        fprintf(stderr, "b2 step trace, stage %d, ring steps of CTA 0 (SM cycles after the stage began): producer at step, slot free, copies issued | wait begin  data there  step done\n", c->step_fine_op);
        for (int k = 0; k < STEP_FINE_STEPS && f[8 * k + 2]; k++)
            fprintf(stderr, "  step %3d | %8lld %8lld %8lld | %8lld %8lld %8lld\n", k, (long long)f[8 * k + 3] - t0, (long long)f[8 * k + 4] - t0,
                    (long long)f[8 * k + 5] - t0, (long long)f[8 * k] - t0, (long long)f[8 * k + 1] - t0, (long long)f[8 * k + 2] - t0);
    }
}

int build_graph(b2_ctx* c) {
    if (!c->use_graph) return 0;
    for (int attempt = 0; attempt < 2; attempt++) {
        const bool pdl = c->use_pdl && attempt == 0;
        cudaGraph_t graph = nullptr;
        cudaGraphExec_t exec = nullptr;
        cudaError_t e = cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal);
        int rc = 0;
        if (e == cudaSuccess) {
            rc = enqueue_decode(c, c->stream, pdl);
            e = cudaStreamEndCapture(c->stream, &graph);
        }
        if (e == cudaSuccess && rc == 0) e = cudaGraphInstantiate(&exec, graph, 0);
        if (e == cudaSuccess && rc == 0) {
            c->graph = graph;
            c->exec = exec;
            c->use_pdl = pdl;
            return 0;
        }
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        if (rc) return rc;  // argument errors do not go away without PDL
        if (!pdl) {
            set_error("ctx: CUDA graph capture failed: %s", cudaGetErrorString(e));
            return B2_ECUDA;
        }
        fprintf(stderr, "b2: graph capture with programmatic dependent launch failed (%s); retrying without\n", cudaGetErrorString(e));
    }
    return B2_ECUDA;
}

int run_step(b2_ctx* c) {
    if (c->exec) {
        B2_CUDA_CHECK(cudaGraphLaunch(c->exec, c->stream));
        return 0;
    }
    return enqueue_decode(c, c->stream, c->use_pdl);
}

}  // namespace

// ---- batched path: up to GEMM_MAX_COLS tokens per pass over the weights (batched decode, chunked prefill) ----

// every matrix in a type the tensor-core kernel computes?
bool batch_supported(const b2_model* m) {
    if (env_flag("B2_NO_BATCH") || m->tp_size > 1) return false;   // shards decode one sequence, one launch per stage
    auto ok = [](const b2_model::W& w) { return gemm_supports_type(w.type); };
    if (!ok(m->output)) return false;
    for (const auto& L : m->layers) {
        if (!ok(L.wq) || !ok(L.wk) || !ok(L.wv) || !ok(L.wo) || !ok(L.gate) || !ok(L.up) || !ok(L.down)) return false;
        // Q, K, V share one launch and one quantised input: same activation format
        if (gemm_act_for_type(L.wq.type) != gemm_act_for_type(L.wk.type) || gemm_act_for_type(L.wq.type) != gemm_act_for_type(L.wv.type)) return false;
    }
    return true;
}

// the batched attention cuts a column's sequence into shorter pieces than the single-sequence launch: more CTAs per column
int batch_attn_splits(int n_ctx) {
    const int s = (n_ctx + ATTN_BATCH_SPLIT_TOKENS - 1) / ATTN_BATCH_SPLIT_TOKENS;
    return s < 1 ? 1 : (s > ATTN_MAX_SPLITS ? ATTN_MAX_SPLITS : s);
}

// attention of n batch columns (positions / slots in device arrays) over their own sequences: the tensor-core kernel
// (attn_batch.cu) when the context fits its chunking, attn.cu's otherwise (B2_ATTN_BATCH_OLD=1 forces the latter)
int enqueue_batch_attention(b2_ctx* c, const float* q, const __half* kc, const __half* vc, float* out, const int32_t* col_pos,
                            const int32_t* col_slot, int n, cudaStream_t st) {
    const b2_model_info& I = c->m->info;
    const int splits = batch_attn_splits(c->n_ctx);
    static const bool old_kernel = env_flag("B2_ATTN_BATCH_OLD");
    if (!old_kernel && attn_batch_supported(I.head_dim, c->n_ctx, splits))
        return launch_attn_batch(q, kc, vc, out, I.n_head, I.n_kv, I.head_dim, c->kv_tokens, col_pos, col_slot, c->slot_stride, n, splits,
                                 c->bt.attn_scratch, c->bt.cap, st, c->d_kv_table, c->pages_per_seq);
    // (only with the static mapping: slot s is the contiguous window at s * slot_stride)
    return launch_attn_decode(q, kc, vc, out, I.n_head, I.n_kv, I.head_dim, c->kv_tokens, col_pos, 0, splits, c->bt.attn_scratch, st, false, n, col_slot,
                              c->slot_stride, ATTN_BATCH_SPLIT_TOKENS, c->bt.cap);
}

// ---- KV pages (model.h) ----
// make positions [p0, p1] of `slot` addressable; with the static mapping everything already is
int kv_map(b2_ctx* c, int slot, int p0, int p1) {
    if (!c->kv_dynamic || slot == 0) return 0;
    int32_t* row = c->kv_table.data() + (size_t)slot * c->pages_per_seq;
    for (int i = p0 / KV_PAGE; i <= p1 / KV_PAGE; i++) {
        if (row[i] >= 0) continue;
        if (c->kv_free.empty()) {
            set_error("KV pool exhausted: %d pages of %d positions, all mapped (b2_ctx_slot_truncate releases a sequence's pages; "
                      "B2_KV_POOL_TOKENS sizes the pool)", c->n_pages, KV_PAGE);
            return B2_ENOMEM;
        }
        row[i] = c->kv_free.back();
        c->kv_free.pop_back();
        c->kv_table_dirty = true;
    }
    return 0;
}

// Attention walks pages 0 .. pos / KV_PAGE of a slot: on an on-demand pool every one of them must have been mapped by
// an earlier pass (the server fills positions contiguously; a direct caller that jumps to position p of an empty slot
// would make the kernels address page -1, in front of the pool).
int kv_check_prefix(const b2_ctx* c, int slot, int p0) {
    if (!c->kv_dynamic || slot == 0) return 0;
    const int32_t* row = c->kv_table.data() + (size_t)slot * c->pages_per_seq;
    for (int i = 0; i < p0 / KV_PAGE; i++)
        if (row[i] < 0) {
            set_error("slot %d: position %d needs positions %d..%d in the KV cache, but that page was never written (paged KV pool: "
                      "fill a sequence from position 0 upwards)", slot, p0, i * KV_PAGE, (i + 1) * KV_PAGE - 1);
            return B2_ERANGE;
        }
    return 0;
}

// the device copy of the table follows the host's before the next pass is enqueued (stream order)
int kv_sync_table(b2_ctx* c) {
    if (!c->kv_table_dirty) return 0;
    B2_CUDA_CHECK(cudaMemcpyAsync(c->d_kv_table, c->kv_table.data(), c->kv_table.size() * 4, cudaMemcpyHostToDevice, c->stream));
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));   // pageable source: keep the host vector free to change
    c->kv_table_dirty = false;
    return 0;
}

int ensure_batch(b2_ctx* c) {
    if (c->bt.cap) return 0;
    const b2_model_info& I = c->m->info;
    const int cap = BATCH_MAX_COLS;
    int rc;
    auto& b = c->bt;
    if ((rc = dmalloc(&b.x, (size_t)cap * I.d_model))) return rc;
    if ((rc = dmalloc(&b.q, (size_t)cap * I.n_head * I.head_dim))) return rc;
    if ((rc = dmalloc(&b.attn_out, (size_t)cap * I.n_head * I.head_dim))) return rc;
    if ((rc = dmalloc(&b.h, (size_t)cap * I.d_ff))) return rc;
    if ((rc = dmalloc(&b.logits, (size_t)cap * I.vocab))) return rc;
    const int max_k = std::max(std::max(I.d_model, I.d_ff), I.n_head * I.head_dim);
    if ((rc = dmalloc(&b.act, std::max(gemm_act_bytes(max_k, GEMM_MAX_COLS, ACT_Q8K), gemm_act_bytes(max_k, GEMM_MAX_COLS, ACT_Q80))))) return rc;
    if ((rc = dmalloc(&b.tok, cap))) return rc;
    if ((rc = dmalloc(&b.pos, cap))) return rc;
    if ((rc = dmalloc(&b.slot, cap))) return rc;
    if ((rc = dmalloc(&b.out, cap))) return rc;
    const size_t asz = attn_scratch_bytes_batch(I.n_head, I.head_dim, batch_attn_splits(c->n_ctx), cap);
    B2_CUDA_CHECK(cudaMalloc(&b.attn_scratch, asz));
    B2_CUDA_CHECK(cudaMemsetAsync(b.attn_scratch, 0, asz, c->stream));
    B2_CUDA_CHECK(cudaMalloc(&b.argmax_scratch, argmax_scratch_bytes() * cap));
    B2_CUDA_CHECK(cudaMemsetAsync(b.argmax_scratch, 0, argmax_scratch_bytes() * cap, c->stream));
    b.cap = cap;
    return 0;
}

// final norm + lm_head + greedy token of n_logits rows of the residual stream -> c->bt.logits / c->bt.out
int enqueue_logits(b2_ctx* c, const float* x_rows, int n_logits, cudaStream_t st) {
    b2_model* m = c->m;
    const b2_model_info& I = m->info;
    auto& b = c->bt;
    int rc;
    if ((rc = launch_act_quant(x_rows, I.d_model, (const float*)m->output_norm.p, I.eps, I.d_model, n_logits, b.act, st,
                               gemm_act_for_type(m->output.type))))
        return rc;
    GemmArgs a{};
    a.K = I.d_model;
    a.act = b.act;
    a.cols.n = n_logits;
    a.out_stride = I.vocab;
    a.slot_stride = c->slot_stride;
    a.rope_tab = c->rope_tab;
    a.rope_half = I.rope_dim / 2;
    seg_matrix(a.seg[0], m->output, PAIR_ADJ, EPI_STORE);
    a.seg[0].out = b.logits;
    a.n_seg = 1;
    if ((rc = gemm_finalize(a))) return rc;
    if ((rc = launch_gemm(a, st))) return rc;
    for (int j = 0; j < n_logits; j++)
        if ((rc = launch_argmax(b.logits + (size_t)j * I.vocab, I.vocab, b.out + j, (uint8_t*)b.argmax_scratch + (size_t)j * argmax_scratch_bytes(),
                                nullptr, nullptr, nullptr, 0, st, false)))
            return rc;
    return 0;
}

// the tensor-core pass (defined with the prompt path below)
bool prefill_tc_supported(const b2_model* m);
int ensure_prefill(b2_ctx* c);
int tc_min_rows(const b2_model* m);
int tcq_min_rows();
bool tcq_pass_ok(const b2_model* m, int T);
struct PromptSeg { int row0, T, pos0, slot; };   // rows [row0, row0 + T) of a pass = positions [pos0, pos0 + T) of `slot`
int enqueue_tc_pass(b2_ctx* c, const int32_t* tok_dev, int T, int slot, int pos0, const int32_t* col_pos, const int32_t* col_slot,
                    cudaStream_t st, const std::vector<PromptSeg>* segs = nullptr);
int enqueue_logits_tc(b2_ctx* c, const float* x_rows, int n_logits, cudaStream_t st);

// does a decode batch of n columns take the tensor-core pass?  Wider than the int8 kernel's 32 columns: always (the
// caller has checked that the pass exists); otherwise from tc_min_rows() columns when the fp16 copy is resident.
bool batch_uses_tc(const b2_ctx* c, int n) {
    if (n > GEMM_MAX_COLS) return true;
    if (c->pf.cap && n >= tcq_min_rows() && tcq_pass_ok(c->m, n)) return true;   // raw blocks dequantised inside the tcgen05 kernel
    return c->pf.cap && c->m->f16_cache && n >= tc_min_rows(c->m);
}

// one pass over the weights for n columns whose (token, slot, position) are in c->bt.tok/slot/pos.
// logits (and the greedy token per column in c->bt.out) are computed for the last n_logits columns.
int enqueue_batch(b2_ctx* c, int n, int n_logits, cudaStream_t st) {
    b2_model* m = c->m;
    const b2_model_info& I = m->info;
    auto& b = c->bt;
    int rc;
    if (batch_uses_tc(c, n)) {
        if ((rc = enqueue_tc_pass(c, b.tok, n, 0, 0, b.pos, b.slot, st))) return rc;
        if (n_logits <= 0) return 0;
        const float* rows = c->pf.x + (size_t)(n - n_logits) * I.d_model;
        const bool lm_tc = (m->f16_cache && n_logits >= tc_min_rows(m)) ||
                           (n_logits >= tcq_min_rows() && tcq_pass_ok(m, n_logits) && tcq_supported(m->output.type, I.d_model));
        return lm_tc ? enqueue_logits_tc(c, rows, n_logits, st) : enqueue_logits(c, rows, n_logits, st);
    }
    const int64_t kv_layer = (int64_t)I.n_kv * c->kv_tokens * I.head_dim;
    const int rope_pairing = I.arch == 1 ? PAIR_NEOX : PAIR_ADJ;
    const int qdim = I.n_head * I.head_dim;
    for (int j = 0; j < n; j++)
        if ((rc = launch_embed(m->tok_embd.type, m->tok_embd.p, row_bytes(m->tok_embd.type, I.d_model), I.d_model, I.vocab, b.tok + j,
                               b.x + (size_t)j * I.d_model, st, false)))
            return rc;
    auto common = [&](GemmArgs& a, int K, int cols, int64_t out_stride) {
        a = GemmArgs{};
        a.K = K;
        a.act = b.act;
        a.cols.n = cols;
        a.cols.pos = b.pos;
        a.cols.slot = b.slot;
        a.out_stride = out_stride;
        a.slot_stride = c->slot_stride;
        a.kv_table = c->d_kv_table;
        a.kv_pps = c->pages_per_seq;
        a.rope_tab = c->rope_tab;
        a.rope_half = I.rope_dim / 2;
    };
    auto run = [&](GemmArgs& a) -> int {
        if ((rc = gemm_finalize(a))) return rc;
        return launch_gemm(a, st);
    };
    GemmArgs a;
    for (int il = 0; il < I.n_layer; il++) {
        const auto& L = m->layers[il];
        __half* kc = c->k_cache + il * kv_layer;
        __half* vc = c->v_cache + il * kv_layer;
        // RMSNorm -> Q, K, V (+bias) -> RoPE -> cache
        if ((rc = launch_act_quant(b.x, I.d_model, (const float*)L.attn_norm.p, I.eps, I.d_model, n, b.act, st, gemm_act_for_type(L.wq.type)))) return rc;
        common(a, I.d_model, n, qdim);
        seg_matrix(a.seg[0], L.wq, rope_pairing, EPI_ROPE_Q);
        a.seg[0].out = b.q;
        a.seg[0].bias = L.bq;
        a.seg[0].head_dim = I.head_dim;
        seg_matrix(a.seg[1], L.wk, rope_pairing, EPI_ROPE_K);
        a.seg[1].cache = kc;
        a.seg[1].bias = L.bk;
        a.seg[1].head_dim = I.head_dim;
        a.seg[1].n_ctx = c->kv_tokens;
        seg_matrix(a.seg[2], L.wv, PAIR_ADJ, EPI_V);
        a.seg[2].cache = vc;
        a.seg[2].bias = L.bv;
        a.seg[2].head_dim = I.head_dim;
        a.seg[2].n_ctx = c->kv_tokens;
        a.n_seg = 3;
        if ((rc = run(a))) return rc;
        // attention of every column over its own sequence
        if ((rc = enqueue_batch_attention(c, b.q, kc, vc, b.attn_out, b.pos, b.slot, n, st))) return rc;
        // output projection + residual
        if ((rc = launch_act_quant(b.attn_out, qdim, nullptr, 0.0f, qdim, n, b.act, st, gemm_act_for_type(L.wo.type)))) return rc;
        common(a, qdim, n, I.d_model);
        seg_matrix(a.seg[0], L.wo, PAIR_ADJ, EPI_RESID);
        a.seg[0].out = b.x;
        a.n_seg = 1;
        if ((rc = run(a))) return rc;
        // RMSNorm -> gate, up -> SwiGLU
        if ((rc = launch_act_quant(b.x, I.d_model, (const float*)L.ffn_norm.p, I.eps, I.d_model, n, b.act, st, gemm_act_for_type(L.gate.type)))) return rc;
        if (L.gate.type != L.up.type) {
            set_error("model: ffn_gate and ffn_up of layer %d have different types: unsupported", il);
            return B2_EUNSUP;
        }
        common(a, I.d_model, n, I.d_ff);
        seg_matrix(a.seg[0], L.gate, PAIR_SPLIT, EPI_SWIGLU);
        a.seg[0].wb = L.up.p;
        a.seg[0].out = b.h;
        a.n_seg = 1;
        if ((rc = run(a))) return rc;
        // down projection + residual
        if ((rc = launch_act_quant(b.h, I.d_ff, nullptr, 0.0f, I.d_ff, n, b.act, st, gemm_act_for_type(L.down.type)))) return rc;
        common(a, I.d_ff, n, I.d_model);
        seg_matrix(a.seg[0], L.down, PAIR_ADJ, EPI_RESID);
        a.seg[0].out = b.x;
        a.n_seg = 1;
        if ((rc = run(a))) return rc;
    }
    // final norm + lm_head + greedy token for the last n_logits columns (none for the inner chunks of a prompt)
    if (n_logits <= 0) return 0;
    return enqueue_logits(c, b.x + (size_t)(n - n_logits) * I.d_model, n_logits, st);
}

// host arrays -> device columns (through the pinned staging buffer), then enqueue_batch
int run_batch(b2_ctx* c, const int32_t* tokens, const int32_t* slots, const int32_t* pos, int n, int n_logits) {
    int rc;
    if ((rc = ensure_batch(c))) return rc;
    if (n >= 16 && !c->pf.cap && prefill_tc_supported(c->m) && (rc = ensure_prefill(c))) return rc;
    if (n > GEMM_MAX_COLS && !(c->pf.cap && (c->m->f16_cache || tcq_pass_ok(c->m, n)))) {
        set_error("b2_ctx_decode_batch: %d columns need the tensor-core pass: a resident fp16 copy of the model, or at most 128 columns of Q4_K / "
                  "Q6_K tensors (b2_ctx_max_batch)", n);
        return B2_EUNSUP;
    }
    for (int j = 0; j < n; j++) {
        if ((rc = kv_check_prefix(c, slots ? slots[j] : 0, pos[j]))) return rc;
        if ((rc = kv_map(c, slots ? slots[j] : 0, pos[j], pos[j]))) return rc;
    }
    if ((rc = kv_sync_table(c))) return rc;
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));   // staging buffer reuse
    for (int j = 0; j < n; j++) {
        c->h_pin[j] = tokens[j];
        c->h_pin[BATCH_MAX_COLS + j] = pos[j];
        c->h_pin[2 * BATCH_MAX_COLS + j] = slots ? slots[j] : 0;
    }
    B2_CUDA_CHECK(cudaMemcpyAsync(c->bt.tok, c->h_pin, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    B2_CUDA_CHECK(cudaMemcpyAsync(c->bt.pos, c->h_pin + BATCH_MAX_COLS, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    B2_CUDA_CHECK(cudaMemcpyAsync(c->bt.slot, c->h_pin + 2 * BATCH_MAX_COLS, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    if (!c->use_graph) return enqueue_batch(c, n, n_logits, c->stream);
    auto& graphs = c->bt.graphs;
    if (graphs.empty()) graphs.resize((size_t)2 * (BATCH_MAX_COLS + 1) * (BATCH_MAX_COLS + 1));
    // one graph per (columns, logits rows) and per mat-mul path: B2_TCQ can be switched between calls
    const size_t variant = (c->pf.cap && tcq_pass_ok(c->m, n)) ? 1 : 0;
    auto& g = graphs[(variant * (BATCH_MAX_COLS + 1) + (size_t)n) * (BATCH_MAX_COLS + 1) + n_logits];
    if (g.exec) {
        B2_CUDA_CHECK(cudaGraphLaunch(g.exec, c->stream));
        return 0;
    }
    if (!g.seen) {   // first use of this shape: eager (lazy module loading, function attributes)
        g.seen = true;
        return enqueue_batch(c, n, n_logits, c->stream);
    }
    cudaGraph_t graph = nullptr;
    B2_CUDA_CHECK(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    rc = enqueue_batch(c, n, n_logits, c->stream);
    const cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    if (rc) {
        if (graph) cudaGraphDestroy(graph);
        return rc;
    }
    B2_CUDA_CHECK(e);
    const cudaError_t ei = cudaGraphInstantiate(&g.exec, graph, 0);
    cudaGraphDestroy(graph);
    B2_CUDA_CHECK(ei);
    B2_CUDA_CHECK(cudaGraphLaunch(g.exec, c->stream));
    return 0;
}

// ---- prompt phase on the tcgen05 path: whole chunks of one sequence's prompt per pass over the weights ----

// prompts at least this long take the tensor-core path (below it the 32-column int8 kernel is faster: one pass
// of the tensor-core path dequantises every matrix to fp16 first)
int tc_min_rows(const b2_model* m) {
    static const int v = [] {
        const char* e = getenv("B2_TC_MIN_ROWS");
        return e ? atoi(e) : 0;
    }();
    if (v > 0) return v;
    return m->f16_cache ? 24 : 48;   // with the fp16 copy resident a pass is one HBM sweep: measured faster than the int8 kernel beyond ~20 rows
}

// Can a pass of T rows run every layer mat-mul on the fused-dequantisation kernel?  (Q4_K / Q6_K tensors with row strides a
// tensor map accepts; B2_NO_TCQ=1 switches it off)
bool tcq_model_ok(const b2_model* m) {
    if (env_flag("B2_NO_TCQ")) return false;
    const b2_model_info& I = m->info;
    const int qdim = I.n_head * I.head_dim;
    for (const auto& L : m->layers) {
        if (!tcq_supported(L.wq.type, I.d_model) || !tcq_supported(L.wk.type, I.d_model) || !tcq_supported(L.wv.type, I.d_model) ||
            !tcq_supported(L.wo.type, qdim) || !tcq_supported(L.gate.type, I.d_model) || !tcq_supported(L.up.type, I.d_model) ||
            !tcq_supported(L.down.type, I.d_ff))
            return false;
        if ((L.wq.rows | L.wk.rows | L.wv.rows | L.gate.rows) & 3) return false;   // stacked segments start on multiples of 4 rows
    }
    return true;
}
// Which passes take it: measured on the 8B shape, the fused kernel beats the int8 kernel from ~24 columns (5.06 vs 5.91 ms
// at 32) and is within 4 % of the pass over the resident fp16 copy (5.59 vs 5.35 ms at 64 columns) while reading 3.2 x
// fewer bytes; so by default it serves the models whose fp16 copy is not resident (it does not fit, or B2_NO_F16_CACHE=1)
// -- those used to dequantise every matrix again per pass -- and B2_TCQ=1 makes it the path of every pass of <= 128 rows.
bool tcq_pass_ok(const b2_model* m, int T) {
    const bool force = env_flag("B2_TCQ");   // read per call: the tests switch it inside one process
    return T <= TCQ_MAX_T && tcq_model_ok(m) && (force || !m->f16_cache);
}
// columns from which a decode batch takes the fused tensor-core pass instead of the int8 kernel
int tcq_min_rows() {
    static const int v = [] {
        const char* e = getenv("B2_TCQ_MIN_ROWS");
        return e ? atoi(e) : 0;
    }();
    return v > 0 ? v : 24;
}

int prefill_tc_chunk(const b2_ctx* c) {
    const char* e = getenv("B2_PREFILL_CHUNK");
    int x = e ? atoi(e) : 0;
    if (x <= 0) x = 2048;
    return std::min(x, c->n_ctx);
}

bool prefill_tc_supported(const b2_model* m) {
    if (env_flag("B2_NO_TC_PREFILL") || !batch_supported(m)) return false;
    const b2_model_info& I = m->info;
    if (I.head_dim != 64 && I.head_dim != 128) return false;
    if (I.d_model % TC_BK || I.d_ff % TC_BK || (I.n_head * I.head_dim) % TC_BK) return false;
    for (const auto& L : m->layers) {
        const b2_model::W* ws[] = {&L.wq, &L.wk, &L.wv, &L.wo, &L.gate, &L.up, &L.down};
        for (const b2_model::W* w : ws)
            if (!type_traits(w->type)) return false;
    }
    return true;
}

// fp16 copies of every layer matrix, built once per model when the budget allows (see model.h)
int ensure_f16_cache(b2_model* m, cudaStream_t st) {
    std::lock_guard<std::mutex> lk(m->f16_mu);
    if (m->f16_tried) return 0;
    m->f16_tried = true;
    if (env_flag("B2_NO_F16_CACHE")) return 0;
    const b2_model_info& I = m->info;
    const size_t qdim = (size_t)I.n_head * I.head_dim, kvdim = (size_t)I.n_kv * I.head_dim, d = (size_t)I.d_model, ff = (size_t)I.d_ff;
    const size_t per_layer = (qdim + 2 * kvdim) * d + d * qdim + 2 * ff * d + d * ff;
    const size_t out_elems = (size_t)I.vocab * d;   // lm_head: the logits of wide decode batches
    const size_t bytes = (per_layer * I.n_layer + out_elems) * 2;
    const char* e = getenv("B2_F16_CACHE_GB");
    const double cap_gb = e ? atof(e) : 64.0;
    size_t free_b = 0, total_b = 0;
    B2_CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
    if ((double)bytes > cap_gb * 1e9 || bytes > free_b / 2) return 0;
    void* p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    __half* base = (__half*)p;
    std::vector<b2_model::F16Layer> off(I.n_layer);
    size_t o = 0;
    int rc = 0;
    auto dq = [&](const b2_model::W& w, size_t at) { return launch_dequant_f16(w.type, w.p, base + at, (int64_t)w.rows * w.cols, st); };
    for (int il = 0; il < I.n_layer && !rc; il++) {
        const auto& L = m->layers[il];
        off[il].qkv = o;
        if (!rc) rc = dq(L.wq, o);
        if (!rc) rc = dq(L.wk, o + qdim * d);
        if (!rc) rc = dq(L.wv, o + (qdim + kvdim) * d);
        o += (qdim + 2 * kvdim) * d;
        off[il].o = o;
        if (!rc) rc = dq(L.wo, o);
        o += d * qdim;
        off[il].gate_up = o;
        if (!rc) rc = dq(L.gate, o);
        if (!rc) rc = dq(L.up, o + ff * d);
        o += 2 * ff * d;
        off[il].down = o;
        if (!rc) rc = dq(L.down, o);
        o += d * ff;
    }
    m->f16_output = o;
    if (!rc) rc = dq(m->output, o);
    if (!rc && cudaStreamSynchronize(st) != cudaSuccess) rc = B2_ECUDA;
    if (rc) {
        cudaFree(p);
        return rc;
    }
    m->f16_off = std::move(off);
    m->f16_cache = p;
    return 0;
}

int ensure_prefill(b2_ctx* c) {
    auto& p = c->pf;
    if (p.cap) return 0;
    const b2_model_info& I = c->m->info;
    // rows: a prompt chunk, and never fewer than the widest decode batch (the batch form of the pass uses the same buffers)
    const size_t cap = (size_t)std::max(prefill_tc_chunk(c), BATCH_MAX_COLS);
    const size_t qdim = (size_t)I.n_head * I.head_dim, kvdim = (size_t)I.n_kv * I.head_dim;
    const size_t max_k = std::max(std::max((size_t)I.d_model, (size_t)I.d_ff), qdim);
    const size_t max_w = std::max(std::max((qdim + 2 * kvdim) * I.d_model, (size_t)I.d_model * qdim), 2 * (size_t)I.d_ff * I.d_model);
    int rc;
    if ((rc = ensure_f16_cache(c->m, c->stream))) return rc;
    if ((rc = dmalloc(&p.x, cap * I.d_model))) return rc;
    if ((rc = dmalloc(&p.qkv, cap * (qdim + 2 * kvdim)))) return rc;
    if ((rc = dmalloc(&p.q, cap * qdim))) return rc;
    if ((rc = dmalloc(&p.attn_out, cap * qdim))) return rc;
    if ((rc = dmalloc(&p.gu, cap * 2 * I.d_ff))) return rc;
    B2_CUDA_CHECK(cudaMalloc(&p.xb, cap * max_k * 2));
    if (!c->m->f16_cache) B2_CUDA_CHECK(cudaMalloc(&p.wb, max_w * 2));   // one matrix at a time, dequantised per pass
    if ((rc = dmalloc(&p.tok, cap))) return rc;
    if ((rc = dmalloc(&p.col_pos, cap))) return rc;
    if ((rc = dmalloc(&p.col_slot, cap))) return rc;
    // split-K workspace: the largest any mat-mul of a pass of up to `cap` rows asks for
    const int shapes[5][2] = {{(int)(qdim + 2 * kvdim), I.d_model}, {I.d_model, (int)qdim}, {2 * I.d_ff, I.d_model}, {I.d_model, I.d_ff}, {I.vocab, I.d_model}};
    for (const auto& s : shapes)
        for (size_t T = 128; T < 2 * cap; T *= 2) p.ws_bytes = std::max(p.ws_bytes, tc_gemm_ws_bytes(s[0], s[1], (int)std::min(T, cap)));
    if (tcq_model_ok(c->m))
        for (const auto& s : shapes)
            if (s[0] % 4 == 0) p.ws_bytes = std::max(p.ws_bytes, tcq_ws_bytes(s[0], s[1], TCQ_MAX_T));
    if (p.ws_bytes) B2_CUDA_CHECK(cudaMalloc(&p.ws, p.ws_bytes));
    p.cap = (int)cap;
    return 0;
}

// One pass of T rows over the layers on the tensor-core path; the residual stream of every row ends up in c->pf.x.
//   prompt form (col_pos == nullptr): the rows are consecutive positions [pos0, pos0 + T) of sequence slot `slot`,
//     causal attention over the chunk (prefill_attn.cu)
//   batch form: row j is column j of a decode batch with its own position col_pos[j] and slot col_slot[j] (device
//     arrays), attention per column over its own sequence (attn_batch.cu)
//   several prompts (segs != nullptr): the batch form's per-row positions and slots, with causal attention per segment
int enqueue_tc_pass(b2_ctx* c, const int32_t* tok_dev, int T, int slot, int pos0, const int32_t* col_pos, const int32_t* col_slot,
                    cudaStream_t st, const std::vector<PromptSeg>* segs) {
    b2_model* m = c->m;
    const b2_model_info& I = m->info;
    auto& p = c->pf;
    int rc;
    const int qdim = I.n_head * I.head_dim, kvdim = I.n_kv * I.head_dim;
    const int64_t kv_layer = (int64_t)I.n_kv * c->kv_tokens * I.head_dim;
    __half* xb = (__half*)p.xb;
    __half* wb = (__half*)p.wb;
    auto dq = [&](const b2_model::W& w, __half* dst) { return launch_dequant_f16(w.type, w.p, dst, (int64_t)w.rows * w.cols, st); };
    const __half* cache = (const __half*)m->f16_cache;
    // B2_TC_TRACE=1 (with B2_NO_GRAPH=1): device time per kernel family of this pass, printed to stderr
    static const bool trace = env_flag("B2_TC_TRACE");
    std::vector<cudaEvent_t> ev;
    std::vector<int> ev_cat;
    auto mark = [&](int cat) {   // the work enqueued since the previous mark belongs to family `cat`
        if (!trace) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, st);
        ev.push_back(e);
        ev_cat.push_back(cat);
    };
    mark(-1);
    // defer != nullptr: a split-K reduction (and the epilogue) is left to the kernel that consumes the output
    auto mm = [&](const __half* w, int N, int K, float* out, int64_t ldo, int epi, TcParts* defer) {
        TcGemmArgs a{};
        a.w = w;
        a.x = xb;
        a.out = out;
        a.N = N;
        a.K = K;
        a.T = T;
        a.ldo = ldo;
        a.epi = epi;
        a.ws = p.ws;
        a.ws_bytes = p.ws_bytes;
        mark(1);
        const int r = launch_tc_gemm(a, st, defer);
        mark(0);
        return r;
    };
    // Decode batches (T <= 128 rows): the mat-mul whose producer dequantises the raw blocks (tcq_gemm.cu) -- the quantised
    // tensors cross HBM once, no fp16 copy is read.  The result is always left as slices; a caller that wants it in `out`
    // (defer == nullptr) gets a reduction kernel.
    const bool fused = tcq_pass_ok(m, T);
    const bool perm = fused && tcq_fast_conversion();   // the fused kernel's fp16 conversion wants the activations' K order permuted
    auto mmq = [&](std::initializer_list<const b2_model::W*> ws, int K, float* out, int64_t ldo, int epi, TcParts* defer) {
        TcqGemmArgs a{};
        int N = 0;
        for (const b2_model::W* w : ws) {
            a.seg[a.n_seg++] = TcqSeg{w->p, w->type, (int)w->rows};
            N += (int)w->rows;
        }
        a.x = xb;
        a.x_perm = perm ? 1 : 0;
        a.K = K;
        a.T = T;
        a.ws = p.ws;
        a.ws_bytes = p.ws_bytes;
        mark(1);
        TcParts parts;
        int r = launch_tcq_gemm(a, st, &parts);
        if (r == 0 && defer) *defer = parts;
        if (r == 0 && !defer) r = launch_tcq_reduce(parts, T, N, out, ldo, nullptr, epi, st);
        mark(0);
        return r;
    };
    TcParts pend{nullptr, 1, 0};   // residual mat-mul (O, down) whose slices the next norm kernel adds into x
    if ((rc = launch_embed_rows(m->tok_embd.type, m->tok_embd.p, row_bytes(m->tok_embd.type, I.d_model), I.d_model, I.vocab, tok_dev, T, p.x, st))) return rc;
    for (int il = 0; il < I.n_layer; il++) {
        const auto& L = m->layers[il];
        // the layer's part of the pool; rows are found through the page table (slot = the column's, or `slot`)
        __half* kc = c->k_cache + il * kv_layer;
        __half* vc = c->v_cache + il * kv_layer;
        // RMSNorm -> one mat-mul over the stacked Q | K | V matrices -> bias, RoPE, KV cache
        if ((rc = launch_rows_to_f16(p.x, I.d_model, (const float*)L.attn_norm.p, I.eps, I.d_model, T, gemm_act_for_type(L.wq.type), xb, st, pend, perm)))
            return rc;
        if (!cache && !fused) {
            if ((rc = dq(L.wq, wb))) return rc;
            if ((rc = dq(L.wk, wb + (size_t)qdim * I.d_model))) return rc;
            if ((rc = dq(L.wv, wb + (size_t)(qdim + kvdim) * I.d_model))) return rc;
        }
        TcParts parts;
        if (fused) rc = mmq({&L.wq, &L.wk, &L.wv}, I.d_model, p.qkv, qdim + 2 * kvdim, TC_EPI_STORE, &parts);
        else rc = mm(cache ? cache + m->f16_off[il].qkv : wb, qdim + 2 * kvdim, I.d_model, p.qkv, qdim + 2 * kvdim, TC_EPI_STORE, &parts);
        if (rc) return rc;
        PrefillQkvArgs qa{};
        qa.qkv = p.qkv;
        qa.parts_ws = parts.ws;
        qa.parts_splits = parts.splits;
        qa.parts_slice = parts.slice;
        qa.bq = L.bq;
        qa.bk = L.bk;
        qa.bv = L.bv;
        qa.q = p.q;
        qa.k_cache = kc;
        qa.v_cache = vc;
        qa.rope_tab = c->rope_tab;
        qa.nq = qdim;
        qa.nk = kvdim;
        qa.head_dim = I.head_dim;
        qa.n_ctx = c->kv_tokens;
        qa.kv_table = c->d_kv_table;
        qa.kv_pps = c->pages_per_seq;
        qa.slot = slot;
        qa.pos0 = pos0;
        qa.rope_half = I.rope_dim / 2;
        qa.neox = I.arch == 1;
        qa.col_pos = col_pos;
        qa.col_slot = col_slot;
        qa.slot_stride = c->slot_stride;
        mark(0);
        if ((rc = launch_prefill_qkv_post(qa, T, st))) return rc;
        mark(1);
        if (segs) {
            for (const PromptSeg& sg : *segs)
                if ((rc = launch_prefill_attn(p.q + (size_t)sg.row0 * qdim, kc, vc, p.attn_out + (size_t)sg.row0 * qdim, I.n_head, I.n_kv, I.head_dim,
                                              c->kv_tokens, sg.T, sg.pos0, st, c->d_kv_table + (size_t)sg.slot * c->pages_per_seq)))
                    return rc;
        } else if (col_pos) {
            if ((rc = enqueue_batch_attention(c, p.q, kc, vc, p.attn_out, col_pos, col_slot, T, st))) return rc;
        } else if ((rc = launch_prefill_attn(p.q, kc, vc, p.attn_out, I.n_head, I.n_kv, I.head_dim, c->kv_tokens, T, pos0, st,
                                             c->d_kv_table + (size_t)slot * c->pages_per_seq))) {
            return rc;
        }
        mark(2);
        // output projection + residual
        if ((rc = launch_rows_to_f16(p.attn_out, qdim, nullptr, 0.0f, qdim, T, gemm_act_for_type(L.wo.type), xb, st, TcParts{nullptr, 1, 0}, perm))) return rc;
        if (!cache && !fused && (rc = dq(L.wo, wb))) return rc;
        if (fused) rc = mmq({&L.wo}, qdim, p.x, I.d_model, TC_EPI_ADD, &pend);
        else rc = mm(cache ? cache + m->f16_off[il].o : wb, I.d_model, qdim, p.x, I.d_model, TC_EPI_ADD, &pend);
        if (rc) return rc;
        // (+ the O slices) RMSNorm -> stacked gate | up -> SwiGLU -> down + residual
        if ((rc = launch_rows_to_f16(p.x, I.d_model, (const float*)L.ffn_norm.p, I.eps, I.d_model, T, gemm_act_for_type(L.gate.type), xb, st, pend, perm)))
            return rc;
        if (!cache && !fused) {
            if ((rc = dq(L.gate, wb))) return rc;
            if ((rc = dq(L.up, wb + (size_t)I.d_ff * I.d_model))) return rc;
        }
        if (fused) rc = mmq({&L.gate, &L.up}, I.d_model, p.gu, 2 * (int64_t)I.d_ff, TC_EPI_STORE, &parts);
        else rc = mm(cache ? cache + m->f16_off[il].gate_up : wb, 2 * I.d_ff, I.d_model, p.gu, 2 * (int64_t)I.d_ff, TC_EPI_STORE, &parts);
        if (rc) return rc;
        if ((rc = launch_swiglu_to_f16(p.gu, parts, I.d_ff, T, gemm_act_for_type(L.down.type), xb, il == I.n_layer - 1 ? c->h : nullptr, st, perm))) return rc;
        if (!cache && !fused && (rc = dq(L.down, wb))) return rc;
        // the last layer's down projection reduces its own slices: nothing after it in this pass reads x through a norm kernel
        const bool last = il == I.n_layer - 1;
        if (fused) rc = mmq({&L.down}, I.d_ff, p.x, I.d_model, TC_EPI_ADD, last ? nullptr : &pend);
        else rc = mm(cache ? cache + m->f16_off[il].down : wb, I.d_model, I.d_ff, p.x, I.d_model, TC_EPI_ADD, last ? nullptr : &pend);
        if (rc) return rc;
    }
    if (trace && ev.size() > 1) {
        cudaStreamSynchronize(st);
        float tot[3] = {0, 0, 0};
        for (size_t i = 1; i < ev.size(); i++) {
            float ms = 0;
            cudaEventElapsedTime(&ms, ev[i - 1], ev[i]);
            if (ev_cat[i] >= 0) tot[ev_cat[i]] += ms;
        }
        for (cudaEvent_t e : ev) cudaEventDestroy(e);
        fprintf(stderr, "[b2 tc pass] rows=%d %s: mat-mul (+split-K reduce) %.3f ms, elementwise %.3f ms, attention %.3f ms\n", T,
                col_pos ? "batch" : "prompt", tot[0], tot[1], tot[2]);
    }
    return 0;
}

// a prompt (or a prompt's tail after a reused prefix) of one sequence: chunks of up to prefill_tc_chunk() tokens;
// the last position's logits and greedy token end up in c->bt.logits / c->bt.out[0] when want_logits
int prefill_tc(b2_ctx* c, int slot, const int32_t* tokens, int n, int pos0, bool want_logits) {
    int rc;
    if ((rc = ensure_batch(c))) return rc;
    if ((rc = ensure_prefill(c))) return rc;
    const b2_model_info& I = c->m->info;
    if ((rc = kv_check_prefix(c, slot, pos0))) return rc;
    if ((rc = kv_map(c, slot, pos0, pos0 + n - 1))) return rc;
    if ((rc = kv_sync_table(c))) return rc;
    for (int done = 0; done < n;) {
        const int T = std::min(n - done, std::min(prefill_tc_chunk(c), c->pf.cap));
        B2_CUDA_CHECK(cudaMemcpyAsync(c->pf.tok, tokens + done, (size_t)T * 4, cudaMemcpyHostToDevice, c->stream));
        if ((rc = enqueue_tc_pass(c, c->pf.tok, T, slot, pos0 + done, nullptr, nullptr, c->stream))) return rc;
        done += T;
        if (done == n) {
            const size_t j = (size_t)T - 1, qdim = (size_t)I.n_head * I.head_dim;
            if (want_logits && (rc = enqueue_logits(c, c->pf.x + j * I.d_model, 1, c->stream))) return rc;
            // the last position's activations where the per-token path keeps them (b2_ctx_get_activation)
            B2_CUDA_CHECK(cudaMemcpyAsync(c->x, c->pf.x + j * I.d_model, (size_t)I.d_model * 4, cudaMemcpyDeviceToDevice, c->stream));
            B2_CUDA_CHECK(cudaMemcpyAsync(c->q, c->pf.q + j * qdim, qdim * 4, cudaMemcpyDeviceToDevice, c->stream));
            B2_CUDA_CHECK(cudaMemcpyAsync(c->attn_out, c->pf.attn_out + j * qdim, qdim * 4, cudaMemcpyDeviceToDevice, c->stream));
        }
        // the token ids are pageable host memory: the copy above has completed when cudaMemcpyAsync returns
    }
    return 0;
}

// final norm + lm_head (fp16 copy) + greedy token of n_logits rows -> c->bt.logits / c->bt.out
int enqueue_logits_tc(b2_ctx* c, const float* x_rows, int n_logits, cudaStream_t st) {
    b2_model* m = c->m;
    const b2_model_info& I = m->info;
    int rc;
    __half* xb = (__half*)c->pf.xb;
    const bool no_tcq = env_flag("B2_NO_TCQ") || (m->f16_cache && !env_flag("B2_TCQ"));
    const bool fused = !no_tcq && n_logits <= TCQ_MAX_T && tcq_supported(m->output.type, I.d_model) && I.vocab % 4 == 0 &&
                       tcq_ws_bytes(I.vocab, I.d_model, n_logits) <= c->pf.ws_bytes;
    const bool perm = fused && tcq_fast_conversion();
    if ((rc = launch_rows_to_f16(const_cast<float*>(x_rows), I.d_model, (const float*)m->output_norm.p, I.eps, I.d_model, n_logits,
                                 gemm_act_for_type(m->output.type), xb, st, TcParts{nullptr, 1, 0}, perm)))
        return rc;
    if (fused) {
        // the lm_head's quantised rows dequantised inside the tcgen05 kernel (tcq_gemm.cu)
        TcqGemmArgs q{};
        q.seg[0] = TcqSeg{m->output.p, m->output.type, I.vocab};
        q.n_seg = 1;
        q.x = xb;
        q.x_perm = perm ? 1 : 0;
        q.K = I.d_model;
        q.T = n_logits;
        q.ws = c->pf.ws;
        q.ws_bytes = c->pf.ws_bytes;
        TcParts parts;
        if ((rc = launch_tcq_gemm(q, st, &parts))) return rc;
        if ((rc = launch_tcq_reduce(parts, n_logits, I.vocab, c->bt.logits, I.vocab, nullptr, TC_EPI_STORE, st))) return rc;
        return launch_argmax_rows(c->bt.logits, I.vocab, n_logits, c->bt.out, st);
    }
    if (!m->f16_cache) return enqueue_logits(c, x_rows, n_logits, st);   // neither path applies: the int8 kernel
    TcGemmArgs a{};
    a.w = (const __half*)m->f16_cache + m->f16_output;
    a.x = xb;
    a.out = c->bt.logits;
    a.N = I.vocab;
    a.K = I.d_model;
    a.T = n_logits;
    a.ldo = I.vocab;
    a.epi = TC_EPI_STORE;
    if ((rc = launch_tc_gemm(a, st))) return rc;
    return launch_argmax_rows(c->bt.logits, I.vocab, n_logits, c->bt.out, st);
}

extern "C" int b2_ctx_max_batch(b2_ctx* c) {
    if (!c) return B2_EINVAL;
    if (!batch_supported(c->m)) return 0;
    if (cudaSetDevice(c->m->device) != cudaSuccess) return GEMM_MAX_COLS;
    if (prefill_tc_supported(c->m) && !ensure_batch(c) && !ensure_prefill(c)) {
        if (c->m->f16_cache) return BATCH_MAX_COLS;
        if (tcq_pass_ok(c->m, TCQ_MAX_T)) return TCQ_MAX_T;   // no fp16 copy: the fused-dequantisation pass, up to 128 columns
    }
    return GEMM_MAX_COLS;
}

// The prompts (or prompt tails) of several sequences in ONE pass over the weights: n_seq segments, segment i = lens[i]
// tokens at positions pos0[i].. of slot slots[i], their token ids concatenated in `tokens`.
extern "C" int b2_ctx_prefill_multi(b2_ctx* c, int n_seq, const int32_t* slots, const int32_t* pos0, const int32_t* lens, const int32_t* tokens,
                                    int32_t* next_tokens_out, float* logits_out_host) {
    if (!c || n_seq < 1 || n_seq > BATCH_MAX_COLS || !slots || !pos0 || !lens || !tokens) {
        set_error("b2_ctx_prefill_multi: bad arguments (%d sequences, at most %d)", n_seq, BATCH_MAX_COLS);
        return B2_EINVAL;
    }
    if (!prefill_tc_supported(c->m)) {
        set_error("b2_ctx_prefill_multi: this model's shapes or types are outside the tensor-core prompt path");
        return B2_EUNSUP;
    }
    const b2_model_info& I = c->m->info;
    B2_CUDA_CHECK(cudaSetDevice(c->m->device));
    int rc;
    if ((rc = ensure_batch(c))) return rc;
    if ((rc = ensure_prefill(c))) return rc;
    std::vector<PromptSeg> segs(n_seq);
    int rows = 0;
    for (int i = 0; i < n_seq; i++) {
        if (slots[i] < 0 || slots[i] >= c->n_slots || lens[i] < 1 || pos0[i] < 0 || (int64_t)pos0[i] + lens[i] > c->n_ctx) {
            set_error("b2_ctx_prefill_multi: segment %d (slot %d, positions %d..+%d) out of range", i, slots[i], pos0[i], lens[i]);
            return B2_ERANGE;
        }
        for (int k = 0; k < i; k++)
            if (slots[k] == slots[i]) {
                set_error("b2_ctx_prefill_multi: slot %d appears twice", slots[i]);
                return B2_EINVAL;
            }
        segs[i] = PromptSeg{rows, lens[i], pos0[i], slots[i]};
        rows += lens[i];
    }
    if (rows > c->pf.cap) {
        set_error("b2_ctx_prefill_multi: %d rows in one pass, at most %d", rows, c->pf.cap);
        return B2_ERANGE;
    }
    for (int r = 0; r < rows; r++)
        if (tokens[r] < 0 || tokens[r] >= I.vocab) {
            set_error("b2_ctx_prefill_multi: token %d at row %d is outside the vocabulary (%d)", tokens[r], r, I.vocab);
            return B2_ERANGE;
        }
    for (const PromptSeg& sg : segs) {
        if ((rc = kv_check_prefix(c, sg.slot, sg.pos0))) return rc;
        if ((rc = kv_map(c, sg.slot, sg.pos0, sg.pos0 + sg.T - 1))) return rc;
    }
    if ((rc = kv_sync_table(c))) return rc;
    std::vector<int32_t> rpos(rows), rslot(rows);
    for (const PromptSeg& sg : segs)
        for (int t = 0; t < sg.T; t++) {
            rpos[sg.row0 + t] = sg.pos0 + t;
            rslot[sg.row0 + t] = sg.slot;
        }
    B2_CUDA_CHECK(cudaMemcpyAsync(c->pf.tok, tokens, (size_t)rows * 4, cudaMemcpyHostToDevice, c->stream));
    B2_CUDA_CHECK(cudaMemcpyAsync(c->pf.col_pos, rpos.data(), (size_t)rows * 4, cudaMemcpyHostToDevice, c->stream));
    B2_CUDA_CHECK(cudaMemcpyAsync(c->pf.col_slot, rslot.data(), (size_t)rows * 4, cudaMemcpyHostToDevice, c->stream));
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));   // pageable sources
    if ((rc = enqueue_tc_pass(c, c->pf.tok, rows, 0, 0, c->pf.col_pos, c->pf.col_slot, c->stream, &segs))) return rc;
    if (next_tokens_out || logits_out_host) {
        // the last row of every segment, gathered (the Q|K|V buffer is free by now), through the lm_head
        float* gathered = c->pf.qkv;
        for (int i = 0; i < n_seq; i++)
            B2_CUDA_CHECK(cudaMemcpyAsync(gathered + (size_t)i * I.d_model, c->pf.x + (size_t)(segs[i].row0 + segs[i].T - 1) * I.d_model,
                                          (size_t)I.d_model * 4, cudaMemcpyDeviceToDevice, c->stream));
        if (n_seq >= tc_min_rows(c->m) && c->m->f16_cache) rc = enqueue_logits_tc(c, gathered, n_seq, c->stream);
        else if (n_seq <= GEMM_MAX_COLS) rc = enqueue_logits(c, gathered, n_seq, c->stream);
        else {
            set_error("b2_ctx_prefill_multi: %d sequences need the resident fp16 copy for their logits", n_seq);
            rc = B2_EUNSUP;
        }
        if (rc) return rc;
        {
            std::vector<int32_t> ls(n_seq), lp(n_seq);
            for (int i = 0; i < n_seq; i++) { ls[i] = segs[i].slot; lp[i] = segs[i].pos0 + segs[i].T - 1; }
            if ((rc = sample_columns(c, ls.data(), lp.data(), 0, n_seq))) return rc;
        }
        if (logits_out_host)
            B2_CUDA_CHECK(cudaMemcpyAsync(logits_out_host, c->bt.logits, (size_t)n_seq * I.vocab * 4, cudaMemcpyDeviceToHost, c->stream));
        if (next_tokens_out) B2_CUDA_CHECK(cudaMemcpyAsync(c->h_pin, c->bt.out, (size_t)n_seq * 4, cudaMemcpyDeviceToHost, c->stream));
    }
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    if (next_tokens_out) memcpy(next_tokens_out, c->h_pin, (size_t)n_seq * 4);
    return 0;
}

extern "C" int b2_ctx_slot_truncate(b2_ctx* c, int slot, int n_keep) {
    if (!c || slot < 0 || slot >= c->n_slots || n_keep < 0) {
        set_error("b2_ctx_slot_truncate: bad arguments (slot %d, keep %d)", slot, n_keep);
        return B2_EINVAL;
    }
    if (!c->kv_dynamic || slot == 0) return 0;   // statically mapped: nothing to give back
    // the pages may still be read by work in flight
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    int32_t* row = c->kv_table.data() + (size_t)slot * c->pages_per_seq;
    for (int i = (n_keep + KV_PAGE - 1) / KV_PAGE; i < c->pages_per_seq; i++)
        if (row[i] >= 0) {
            c->kv_free.push_back(row[i]);
            row[i] = -1;
            c->kv_table_dirty = true;
        }
    return 0;
}

extern "C" int b2_ctx_kv_pages(const b2_ctx* c, int* total_out, int* free_out, int* page_tokens_out) {
    if (!c) return B2_EINVAL;
    if (total_out) *total_out = c->n_pages;
    if (free_out) *free_out = c->kv_dynamic ? (int)c->kv_free.size() : 0;
    if (page_tokens_out) *page_tokens_out = KV_PAGE;
    return c->kv_dynamic ? 1 : 0;
}

extern "C" int b2_ctx_prefill(b2_ctx* c, int slot, const int32_t* tokens, int n, int pos0, int32_t* next_token_out, float* logits_out_host) {
    if (!c || !tokens || n < 1) {
        set_error("b2_ctx_prefill: bad arguments");
        return B2_EINVAL;
    }
    const b2_model_info& I = c->m->info;
    if (slot < 0 || slot >= c->n_slots || pos0 < 0 || (int64_t)pos0 + n > c->n_ctx) {
        set_error("b2_ctx_prefill: slot %d / positions %d..%d out of range (%d slots, context %d)", slot, pos0, pos0 + n - 1, c->n_slots, c->n_ctx);
        return B2_ERANGE;
    }
    for (int i = 0; i < n; i++)
        if (tokens[i] < 0 || tokens[i] >= I.vocab) {
            set_error("b2_ctx_prefill: token %d at index %d is outside the vocabulary (%d)", tokens[i], i, I.vocab);
            return B2_ERANGE;
        }
    if (!prefill_tc_supported(c->m)) {
        set_error("b2_ctx_prefill: this model's shapes or types are outside the tensor-core prompt path; use b2_ctx_decode_batch");
        return B2_EUNSUP;
    }
    B2_CUDA_CHECK(cudaSetDevice(c->m->device));
    int rc;
    const bool want = next_token_out || logits_out_host;
    if ((rc = prefill_tc(c, slot, tokens, n, pos0, want))) return rc;
    if (want) {
        const int32_t s1 = slot, p1 = pos0 + n - 1;
        if ((rc = sample_columns(c, &s1, &p1, 0, 1))) return rc;
    }
    if (logits_out_host) B2_CUDA_CHECK(cudaMemcpyAsync(logits_out_host, c->bt.logits, (size_t)I.vocab * 4, cudaMemcpyDeviceToHost, c->stream));
    if (next_token_out) B2_CUDA_CHECK(cudaMemcpyAsync(c->h_pin, c->bt.out, 4, cudaMemcpyDeviceToHost, c->stream));
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    if (next_token_out) *next_token_out = c->h_pin[0];
    return 0;
}

extern "C" int b2_ctx_n_slots(const b2_ctx* c) { return c ? c->n_slots : B2_EINVAL; }

extern "C" int b2_ctx_set_sampler(b2_ctx* c, int slot, const b2_sampling* p) {
    if (!c || slot < 0 || slot >= c->n_slots) {
        set_error("b2_ctx_set_sampler: slot %d out of range", slot);
        return B2_EINVAL;
    }
    b2_sampling s = p ? *p : b2_sampling{0.0f, 1.0f, 0.0f, 1, 0};
    if (p && (!(s.top_p > 0.0f) || s.top_p > 1.0f || s.min_p < 0.0f || s.min_p > 1.0f || s.temperature != s.temperature)) {
        set_error("b2_ctx_set_sampler: top_p %g must be in (0, 1], min_p %g in [0, 1]", (double)s.top_p, (double)s.min_p);
        return B2_EINVAL;
    }
    B2_CUDA_CHECK(cudaSetDevice(c->m->device));
    if (c->samplers.empty()) {
        c->samplers.assign(c->n_slots, b2_sampling{0.0f, 1.0f, 0.0f, 1, 0});
        B2_CUDA_CHECK(cudaMalloc((void**)&c->d_samplers, (size_t)c->n_slots * sizeof(b2_sampling)));
        B2_CUDA_CHECK(cudaMalloc((void**)&c->d_col_params, BATCH_MAX_COLS * sizeof(b2_sampling)));
        B2_CUDA_CHECK(cudaMalloc((void**)&c->d_col_draw, BATCH_MAX_COLS * sizeof(uint64_t)));
        B2_CUDA_CHECK(cudaMallocHost((void**)&c->h_col_params, BATCH_MAX_COLS * sizeof(b2_sampling)));
        B2_CUDA_CHECK(cudaMallocHost((void**)&c->h_col_draw, BATCH_MAX_COLS * sizeof(uint64_t)));
    }
    const bool was0 = slot_samples(c, 0);
    c->samplers[slot] = s;
    c->n_samplers = 0;
    for (const auto& x : c->samplers) c->n_samplers += x.temperature > 0.0f && x.top_k != 1;
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    B2_CUDA_CHECK(cudaMemcpy(c->d_samplers, c->samplers.data(), (size_t)c->n_slots * sizeof(b2_sampling), cudaMemcpyHostToDevice));
    // the single-sequence step's CUDA graph carries (or not) the sampling kernel: rebuild it when slot 0 changes sides
    if (slot == 0 && was0 != slot_samples(c, 0) && c->exec) {
        cudaGraphExecDestroy(c->exec);
        cudaGraphDestroy(c->graph);
        c->exec = nullptr;
        c->graph = nullptr;
        int rc;
        if ((rc = build_graph(c))) return rc;
    }
    return 0;
}

extern "C" int b2_ctx_decode_batch(b2_ctx* c, const int32_t* tokens, const int32_t* slots, const int32_t* pos, int n, int last_only,
                                   int32_t* next_tokens_out, float* logits_out_host) {
    if (!c || !tokens || !pos || n < 1 || n > BATCH_MAX_COLS) {
        set_error("b2_ctx_decode_batch: bad arguments (n=%d, at most %d columns per call)", n, BATCH_MAX_COLS);
        return B2_EINVAL;
    }
    const b2_model_info& I = c->m->info;
    if (!batch_supported(c->m)) {
        set_error("b2_ctx_decode_batch: this model has matrices outside Q4_K / Q5_K / Q6_K / Q8_0; the batched kernel does not compute them yet");
        return B2_EUNSUP;
    }
    for (int j = 0; j < n; j++) {
        const int s = slots ? slots[j] : 0;
        if (tokens[j] < 0 || tokens[j] >= I.vocab || pos[j] < 0 || pos[j] >= c->n_ctx || s < 0 || s >= c->n_slots) {
            set_error("b2_ctx_decode_batch: column %d (token %d, slot %d, position %d) out of range", j, tokens[j], s, pos[j]);
            return B2_ERANGE;
        }
    }
    B2_CUDA_CHECK(cudaSetDevice(c->m->device));
    const int n_logits = last_only ? 1 : n;
    int rc;
    if ((rc = run_batch(c, tokens, slots, pos, n, n_logits))) return rc;
    if ((rc = sample_columns(c, slots, pos, n - n_logits, n_logits))) return rc;
    if (logits_out_host)
        B2_CUDA_CHECK(cudaMemcpyAsync(logits_out_host, c->bt.logits, (size_t)n_logits * I.vocab * 4, cudaMemcpyDeviceToHost, c->stream));
    if (next_tokens_out) B2_CUDA_CHECK(cudaMemcpyAsync(c->h_pin, c->bt.out, (size_t)n_logits * 4, cudaMemcpyDeviceToHost, c->stream));
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    if (next_tokens_out) memcpy(next_tokens_out, c->h_pin, (size_t)n_logits * 4);
    return 0;
}

extern "C" int b2_ctx_batch_logits(b2_ctx* c, int col, float* logits_out_host) {
    if (!c || !logits_out_host || !c->bt.cap || col < 0 || col >= c->bt.cap) {
        set_error("b2_ctx_batch_logits: bad arguments (column %d)", col);
        return B2_EINVAL;
    }
    const b2_model_info& I = c->m->info;
    B2_CUDA_CHECK(cudaMemcpyAsync(logits_out_host, c->bt.logits + (size_t)col * I.vocab, (size_t)I.vocab * 4, cudaMemcpyDeviceToHost, c->stream));
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int b2_ctx_create(b2_model* m, int n_ctx, b2_ctx** out) { return b2_ctx_create_slots(m, n_ctx, 1, out); }

extern "C" int b2_ctx_create_slots(b2_model* m, int n_ctx, int n_slots, b2_ctx** out) {
    if (!m || !out) {
        set_error("b2_ctx_create: NULL argument");
        return B2_EINVAL;
    }
    *out = nullptr;
    if (n_slots < 1 || n_slots > 256) {
        set_error("b2_ctx_create_slots: n_slots=%d outside [1, 256]", n_slots);
        return B2_EINVAL;
    }
    const b2_model_info& I = m->info;
    if (n_ctx <= 0) {
        const char* e = getenv("B2_CTX_SIZE");
        n_ctx = e ? atoi(e) : 0;
        if (n_ctx <= 0) n_ctx = I.n_ctx_train < 8192 ? I.n_ctx_train : 8192;
    }
    if (n_ctx > I.n_ctx_train && I.n_ctx_train > 0 && !env_flag("B2_ALLOW_CTX_EXTEND")) n_ctx = I.n_ctx_train;
    B2_CUDA_CHECK(cudaSetDevice(m->device));
    std::unique_ptr<b2_ctx> c(new b2_ctx());
    c->m = m;
    c->n_ctx = n_ctx;
    c->use_graph = !env_flag("B2_NO_GRAPH");
    c->use_pdl = !env_flag("B2_NO_PDL");
    B2_CUDA_CHECK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    int rc;
    // the KV pool (model.h): [layer][kv head][kv_tokens][head_dim], pages of KV_PAGE positions, slot 0 = pages 0..
    c->n_slots = n_slots;
    c->pages_per_seq = (n_ctx + KV_PAGE - 1) / KV_PAGE;
    {
        const size_t full_pages = (size_t)c->pages_per_seq * n_slots;
        const size_t page_bytes = 2ull * I.n_layer * I.n_kv * KV_PAGE * I.head_dim * sizeof(__half);   // K and V
        size_t pages = full_pages;
        if (const char* e = getenv("B2_KV_POOL_TOKENS")) {
            const long long t = atoll(e);
            if (t > 0) pages = (size_t)((t + KV_PAGE - 1) / KV_PAGE);
        }
        size_t free_b = 0, total_b = 0;
        B2_CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
        const size_t cap_pages = (size_t)((double)free_b * 0.6 / (double)page_bytes);   // leave room for the fp16 copy and scratch
        if (n_slots > 1 && pages > cap_pages) pages = cap_pages;
        // at least slot 0 whole and one page for every other slot
        pages = std::max(pages, (size_t)c->pages_per_seq + (size_t)(n_slots - 1));
        pages = std::min(pages, full_pages);
        c->n_pages = (int)pages;
        c->kv_dynamic = pages < full_pages;
    }
    if (c->kv_dynamic && !attn_batch_supported(I.head_dim, n_ctx, batch_attn_splits(n_ctx))) {
        set_error("ctx: a KV pool smaller than n_slots x n_ctx needs head_dim 64/128 and n_ctx <= %d", KV_PAGE * ATTN_MAX_SPLITS);
        return B2_EUNSUP;
    }
    c->kv_tokens = c->n_pages * KV_PAGE;
    c->slot_stride = (int64_t)c->pages_per_seq * KV_PAGE * I.head_dim;   // static mapping: slot s = rows [s * pps * KV_PAGE, ..)
    c->kv_table.assign((size_t)n_slots * c->pages_per_seq, -1);
    for (int s = 0; s < n_slots; s++)
        for (int i = 0; i < c->pages_per_seq; i++)
            if (s == 0 || !c->kv_dynamic) c->kv_table[(size_t)s * c->pages_per_seq + i] = s * c->pages_per_seq + i;
    if (c->kv_dynamic)
        for (int pg = c->n_pages - 1; pg >= c->pages_per_seq; pg--) c->kv_free.push_back(pg);
    if ((rc = dmalloc(&c->d_kv_table, c->kv_table.size()))) return rc;
    B2_CUDA_CHECK(cudaMemcpyAsync(c->d_kv_table, c->kv_table.data(), c->kv_table.size() * 4, cudaMemcpyHostToDevice, c->stream));
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    const size_t kv_elems = (size_t)I.n_layer * I.n_kv * c->kv_tokens * I.head_dim;
    if ((rc = dmalloc(&c->k_cache, kv_elems))) return rc;
    if ((rc = dmalloc(&c->v_cache, kv_elems))) return rc;
    B2_CUDA_CHECK(cudaMemsetAsync(c->k_cache, 0, kv_elems * sizeof(__half), c->stream));
    B2_CUDA_CHECK(cudaMemsetAsync(c->v_cache, 0, kv_elems * sizeof(__half), c->stream));
    if ((rc = dmalloc(&c->x, I.d_model))) return rc;
    if ((rc = dmalloc(&c->q, (size_t)I.n_head * I.head_dim))) return rc;
    if ((rc = dmalloc(&c->attn_out, (size_t)I.n_head * I.head_dim))) return rc;
    if ((rc = dmalloc(&c->h, I.d_ff))) return rc;
    if ((rc = dmalloc(&c->logits, I.vocab))) return rc;
    c->out_cap = 4096;
    c->prompt_cap = 4096;
    if ((rc = dmalloc(&c->d_tok, 1))) return rc;
    if ((rc = dmalloc(&c->d_pos, 1))) return rc;
    if ((rc = dmalloc(&c->d_step, 1))) return rc;
    if ((rc = dmalloc(&c->d_out, c->out_cap))) return rc;
    if ((rc = dmalloc(&c->d_prompt, c->prompt_cap))) return rc;
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_tok, 0, 4, c->stream));
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_pos, 0, 4, c->stream));
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_step, 0, 4, c->stream));
    c->h_pin_cap = 4096;
    B2_CUDA_CHECK(cudaMallocHost((void**)&c->h_pin, c->h_pin_cap * sizeof(int32_t)));
    c->max_splits = attn_max_splits(n_ctx);
    const size_t asz = attn_scratch_bytes(I.n_head, I.head_dim, c->max_splits);
    B2_CUDA_CHECK(cudaMalloc(&c->attn_scratch, asz));
    B2_CUDA_CHECK(cudaMemsetAsync(c->attn_scratch, 0, asz, c->stream));
    B2_CUDA_CHECK(cudaMalloc(&c->argmax_scratch, argmax_scratch_bytes()));
    B2_CUDA_CHECK(cudaMemsetAsync(c->argmax_scratch, 0, argmax_scratch_bytes(), c->stream));
    // RoPE table built on the host the way ggml builds its per-position cache (ggml_rope_cache_init):
    // theta_0 = pos, theta_{i+1} = theta_i * base^(-2/n_dims) in fp32, angle = theta / freq_factor,
    // cosf/sinf from the host libm -- the same libm the CPU reference would call.
    {
        const int half = I.rope_dim / 2;
        std::vector<float2> tab((size_t)n_ctx * half);
        const float theta_scale = powf(I.rope_base, -2.0f / (float)I.rope_dim);
        for (int p = 0; p < n_ctx; p++) {
            float theta = (float)p;
            for (int i = 0; i < half; i++) {
                const float ff = m->rope_freqs_host.empty() ? 1.0f : m->rope_freqs_host[i];
                const float t = theta / ff;
                tab[(size_t)p * half + i] = make_float2(cosf(t), sinf(t));
                theta *= theta_scale;
            }
        }
        if ((rc = dmalloc(&c->rope_tab, tab.size()))) return rc;
        B2_CUDA_CHECK(cudaMemcpyAsync(c->rope_tab, tab.data(), tab.size() * sizeof(float2), cudaMemcpyHostToDevice, c->stream));
        B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    }
    // scratch + program of the persistent per-token kernel (B2_NO_STEP=1 keeps one launch per stage)
    c->use_step = !env_flag("B2_NO_STEP");   // (a shard's program is built by b2_allreduce_init: its exchange stages need the peers)
    if (m->tp_size > 1) {
        // the first step (and the CUDA graph of a step) need the communicator: b2_allreduce_init finishes the set-up
        B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
        c->pos = 0;
        *out = c.release();
        return 0;
    }
    if (c->use_step) {
        B2_CUDA_CHECK(cudaMalloc((void**)&c->step_counters, (8 + I.n_head) * sizeof(unsigned int)));
        B2_CUDA_CHECK(cudaMemsetAsync(c->step_counters, 0, (8 + I.n_head) * sizeof(unsigned int), c->stream));
        B2_CUDA_CHECK(cudaMalloc((void**)&c->step_sync, STEP_SYNC_BYTES));
        B2_CUDA_CHECK(cudaMalloc((void**)&c->step_argmax_val, 1024 * sizeof(float)));
        B2_CUDA_CHECK(cudaMalloc((void**)&c->step_argmax_idx, 1024 * sizeof(int)));
        if ((rc = build_step_program(c.get()))) return rc;
    }
    // run one eager step so lazy module loading and attribute setting happen outside capture
    if ((rc = enqueue_decode(c.get(), c->stream, false))) return rc;
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_pos, 0, 4, c->stream));
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_step, 0, 4, c->stream));
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_tok, 0, 4, c->stream));
    if ((rc = build_graph(c.get()))) return rc;
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    c->pos = 0;
    *out = c.release();
    return 0;
}

extern "C" void b2_ctx_free(b2_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->m->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->exec) cudaGraphExecDestroy(c->exec);
    if (c->graph) cudaGraphDestroy(c->graph);
    if (c->d_kv_table) cudaFree(c->d_kv_table);
    if (c->tp_local) tp_peer_release(c->m->tp_rank, c->m->tp_size, c->tp_local, c->tp_peers);
    if (c->tp_epoch) cudaFree(c->tp_epoch);
    if (c->tp_comm) tp_comm_free(c->tp_comm);
    void* ptrs[] = {c->k_cache, c->v_cache, c->x, c->q, c->attn_out, c->h, c->logits, c->d_tok, c->d_pos, c->d_step,
                    c->d_out, c->d_prompt, c->rope_tab, c->attn_scratch, c->argmax_scratch,
                    c->d_ops, c->step_attn_partials, c->step_counters, c->step_sync, c->step_argmax_val, c->step_argmax_idx, c->step_trace,
                    c->d_samplers, c->d_col_params, c->d_col_draw};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    for (auto& g : c->bt.graphs)
        if (g.exec) cudaGraphExecDestroy(g.exec);
    void* bptrs[] = {c->bt.x, c->bt.q, c->bt.attn_out, c->bt.h, c->bt.logits, c->bt.act, c->bt.tok, c->bt.pos, c->bt.slot, c->bt.out,
                     c->bt.attn_scratch, c->bt.argmax_scratch};
    for (void* p : bptrs)
        if (p) cudaFree(p);
    void* pptrs[] = {c->pf.x, c->pf.qkv, c->pf.q, c->pf.attn_out, c->pf.gu, c->pf.xb, c->pf.wb, c->pf.tok, c->pf.ws, c->pf.col_pos, c->pf.col_slot};
    for (void* p : pptrs)
        if (p) cudaFree(p);
    if (c->h_pin) cudaFreeHost(c->h_pin);
    if (c->h_col_params) cudaFreeHost(c->h_col_params);
    if (c->h_col_draw) cudaFreeHost(c->h_col_draw);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" int b2_allreduce_unique_id(void* id_out, size_t cap) { return tp_unique_id(id_out, cap); }

extern "C" int b2_allreduce_init(b2_ctx* c, const void* id, size_t id_bytes) {
    if (!c) return B2_EINVAL;
    if (c->m->tp_size < 2) {
        set_error("b2_allreduce_init: the model was loaded as a full replica (tp_size 1)");
        return B2_EINVAL;
    }
    if (c->tp_comm) return 0;
    B2_CUDA_CHECK(cudaSetDevice(c->m->device));
    int rc;
    if ((rc = tp_comm_init(id, id_bytes, c->m->tp_rank, c->m->tp_size, &c->tp_comm))) return rc;
    // The persistent kernel on a shard: the all-reduce after O and down becomes a stage of the kernel -- every rank's
    // epilogues store their partial rows into every rank's inbox (peer memory over NVLink, plain stores), the ranks
    // exchange sequence numbers, each sums its inbox in rank order.  B2_TP_NCCL=1 (or a shape the kernel does not take,
    // or more than 8 ranks) keeps one launch per stage with ncclAllReduce between them.
    if (c->use_step && !env_flag("B2_TP_NCCL") && c->m->tp_size <= 8) {
        const b2_model_info& I = c->m->info;
        const int tp = c->m->tp_size;
        c->tp_flags_off = ((size_t)2 * tp * I.d_model * sizeof(float) + 255) & ~(size_t)255;
        if ((rc = tp_peer_buffers(c->tp_comm, c->m->tp_rank, tp, c->tp_flags_off + 256, &c->tp_local, c->tp_peers, c->stream))) return rc;
        B2_CUDA_CHECK(cudaMalloc((void**)&c->tp_epoch, 256));
        B2_CUDA_CHECK(cudaMemset(c->tp_epoch, 0, 256));
        B2_CUDA_CHECK(cudaMalloc((void**)&c->step_counters, (8 + I.n_head) * sizeof(unsigned int)));
        B2_CUDA_CHECK(cudaMemsetAsync(c->step_counters, 0, (8 + I.n_head) * sizeof(unsigned int), c->stream));
        B2_CUDA_CHECK(cudaMalloc((void**)&c->step_sync, STEP_SYNC_BYTES));
        B2_CUDA_CHECK(cudaMalloc((void**)&c->step_argmax_val, 1024 * sizeof(float)));
        B2_CUDA_CHECK(cudaMalloc((void**)&c->step_argmax_idx, 1024 * sizeof(int)));
        if ((rc = build_step_program(c))) return rc;
        // nobody launches a step before every rank has mapped the buffers and cleared its flags
        if ((rc = tp_barrier(c->tp_comm, (float*)((uint8_t*)c->tp_epoch + 128), c->stream))) return rc;
    } else {
        c->use_step = false;
    }
    // what b2_ctx_create does for a replica: one eager step (lazy module loading outside capture), then the graph.
    // Every rank runs the same steps, so the collectives inside match up.
    if ((rc = enqueue_decode(c, c->stream, false))) return rc;
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_pos, 0, 4, c->stream));
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_step, 0, 4, c->stream));
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_tok, 0, 4, c->stream));
    if (!env_flag("B2_TP_NO_GRAPH") && (rc = build_graph(c))) return rc;   // the all-reduces are captured with the step
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    c->pos = 0;
    return 0;
}

extern "C" int b2_ctx_n_ctx(const b2_ctx* c) { return c ? c->n_ctx : B2_EINVAL; }
extern "C" int b2_ctx_pos(const b2_ctx* c) { return c ? c->pos : B2_EINVAL; }
extern "C" int b2_ctx_launches_per_step(const b2_ctx* c) { return c ? c->launches_per_step : B2_EINVAL; }
extern "C" void* b2_ctx_stream(b2_ctx* c) { return c ? (void*)c->stream : nullptr; }

extern "C" int b2_ctx_sync(b2_ctx* c) {
    if (!c) return B2_EINVAL;
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    return 0;
}

extern "C" int b2_ctx_reset(b2_ctx* c) {
    if (!c) return B2_EINVAL;
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_pos, 0, 4, c->stream));
    c->pos = 0;
    return 0;
}

extern "C" int b2_ctx_decode(b2_ctx* c, const int32_t* tokens, int n, int pos0, float* logits_out_host, int32_t* next_token_out) {
    if (!c || !tokens || n <= 0) {
        set_error("b2_ctx_decode: bad arguments");
        return B2_EINVAL;
    }
    const b2_model_info& I = c->m->info;
    if (pos0 < 0 || (int64_t)pos0 + n > c->n_ctx) {
        set_error("b2_ctx_decode: positions %d..%d exceed the context size %d", pos0, pos0 + n - 1, c->n_ctx);
        return B2_ERANGE;
    }
    for (int i = 0; i < n; i++)
        if (tokens[i] < 0 || tokens[i] >= I.vocab) {
            set_error("b2_ctx_decode: token %d at index %d is outside the vocabulary (%d)", tokens[i], i, I.vocab);
            return B2_ERANGE;
        }
    B2_CUDA_CHECK(cudaSetDevice(c->m->device));
    int rc;
    // prompts: chunks of up to 32 tokens per pass over the weights on the tensor-core path (sequence slot 0), the
    // last chunk producing the logits / greedy token of the final position; then the device-resident state
    // (token, position, logits) is left exactly as the per-token path leaves it
    const int n_in = n;
    const bool tc = n >= 16 && prefill_tc_supported(c->m);
    if (tc && (rc = ensure_prefill(c))) return rc;   // builds the fp16 copy when it fits: decides the threshold below
    if (tc && n >= tc_min_rows(c->m)) {
        // long prompts: whole chunks on the tcgen05 path (fp16 tiles of the dequantised matrices)
        if ((rc = prefill_tc(c, 0, tokens, n, pos0, true))) return rc;
        B2_CUDA_CHECK(cudaMemcpyAsync(c->logits, c->bt.logits, (size_t)I.vocab * 4, cudaMemcpyDeviceToDevice, c->stream));
        B2_CUDA_CHECK(cudaMemcpyAsync(c->d_tok, c->bt.out, 4, cudaMemcpyDeviceToDevice, c->stream));
        B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));   // staging buffer reuse
        c->h_pin[0] = pos0 + n;
        B2_CUDA_CHECK(cudaMemcpyAsync(c->d_pos, c->h_pin, 4, cudaMemcpyHostToDevice, c->stream));
        if ((rc = sample_single(c, c->stream, false))) return rc;
        n = 0;
    } else if (n >= 4 && batch_supported(c->m)) {
        int32_t pos[GEMM_MAX_COLS];
        for (int done = 0; done < n;) {
            const int chunk = std::min(n - done, GEMM_MAX_COLS);
            for (int j = 0; j < chunk; j++) pos[j] = pos0 + done + j;
            const bool last = done + chunk == n;
            if ((rc = run_batch(c, tokens + done, nullptr, pos, chunk, last ? 1 : 0))) return rc;
            if (last) {
                const size_t j = (size_t)chunk - 1, qdim = (size_t)I.n_head * I.head_dim;
                B2_CUDA_CHECK(cudaMemcpyAsync(c->logits, c->bt.logits, (size_t)I.vocab * 4, cudaMemcpyDeviceToDevice, c->stream));
                B2_CUDA_CHECK(cudaMemcpyAsync(c->d_tok, c->bt.out, 4, cudaMemcpyDeviceToDevice, c->stream));
                // the last position's activations, where the per-token path keeps them (b2_ctx_get_activation)
                B2_CUDA_CHECK(cudaMemcpyAsync(c->x, c->bt.x + j * I.d_model, (size_t)I.d_model * 4, cudaMemcpyDeviceToDevice, c->stream));
                B2_CUDA_CHECK(cudaMemcpyAsync(c->q, c->bt.q + j * qdim, qdim * 4, cudaMemcpyDeviceToDevice, c->stream));
                B2_CUDA_CHECK(cudaMemcpyAsync(c->attn_out, c->bt.attn_out + j * qdim, qdim * 4, cudaMemcpyDeviceToDevice, c->stream));
                B2_CUDA_CHECK(cudaMemcpyAsync(c->h, c->bt.h + j * I.d_ff, (size_t)I.d_ff * 4, cudaMemcpyDeviceToDevice, c->stream));
            }
            done += chunk;
        }
        B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));   // staging buffer reuse
        c->h_pin[0] = pos0 + n;
        B2_CUDA_CHECK(cudaMemcpyAsync(c->d_pos, c->h_pin, 4, cudaMemcpyHostToDevice, c->stream));
        if ((rc = sample_single(c, c->stream, false))) return rc;
        n = 0;   // nothing left for the per-token loop
    }
    for (int done = 0; done < n;) {
        const int chunk = std::min(n - done, std::min(c->prompt_cap, c->h_pin_cap - 2));
        // pinned staging: [0] = position, [1..] = token ids
        B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));  // staging buffer reuse
        c->h_pin[0] = pos0 + done;
        memcpy(c->h_pin + 1, tokens + done, (size_t)chunk * sizeof(int32_t));
        B2_CUDA_CHECK(cudaMemcpyAsync(c->d_pos, c->h_pin, 4, cudaMemcpyHostToDevice, c->stream));
        B2_CUDA_CHECK(cudaMemcpyAsync(c->d_prompt, c->h_pin + 1, (size_t)chunk * 4, cudaMemcpyHostToDevice, c->stream));
        for (int i = 0; i < chunk; i++) {
            B2_CUDA_CHECK(cudaMemcpyAsync(c->d_tok, c->d_prompt + i, 4, cudaMemcpyDeviceToDevice, c->stream));
            if ((rc = run_step(c))) return rc;
        }
        done += chunk;
    }
    c->pos = pos0 + n_in;
    if (logits_out_host)
        B2_CUDA_CHECK(cudaMemcpyAsync(logits_out_host, c->logits, (size_t)I.vocab * 4, cudaMemcpyDeviceToHost, c->stream));
    if (next_token_out) {
        B2_CUDA_CHECK(cudaMemcpyAsync(c->h_pin, c->d_tok, 4, cudaMemcpyDeviceToHost, c->stream));
        B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
        *next_token_out = c->h_pin[0];
    } else {
        B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    }
    return 0;
}

extern "C" int b2_ctx_enqueue_steps(b2_ctx* c, int n_steps) {
    if (!c || n_steps < 0) return B2_EINVAL;
    if ((int64_t)c->pos + n_steps > c->n_ctx) {
        set_error("b2_ctx_enqueue_steps: %d steps from position %d exceed the context size %d", n_steps, c->pos, c->n_ctx);
        return B2_ERANGE;
    }
    B2_CUDA_CHECK(cudaSetDevice(c->m->device));
    int rc;
    for (int i = 0; i < n_steps; i++)
        if ((rc = run_step(c))) return rc;
    c->pos += n_steps;
    return 0;
}

extern "C" int b2_ctx_generate_greedy(b2_ctx* c, int n_steps, int32_t* tokens_out_host) {
    if (!c || n_steps < 0 || (n_steps && !tokens_out_host)) return B2_EINVAL;
    int rc;
    for (int done = 0; done < n_steps;) {
        const int chunk = std::min(n_steps - done, std::min(c->out_cap, c->h_pin_cap));
        B2_CUDA_CHECK(cudaMemsetAsync(c->d_step, 0, 4, c->stream));
        if ((rc = b2_ctx_enqueue_steps(c, chunk))) return rc;
        B2_CUDA_CHECK(cudaMemcpyAsync(c->h_pin, c->d_out, (size_t)chunk * 4, cudaMemcpyDeviceToHost, c->stream));
        B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
        memcpy(tokens_out_host + done, c->h_pin, (size_t)chunk * 4);
        done += chunk;
    }
    return 0;
}

extern "C" int b2_ctx_get_activation(b2_ctx* c, int which, float* out_host, int64_t cap) {
    if (!c || !out_host) return B2_EINVAL;
    const b2_model_info& I = c->m->info;
    const float* src = nullptr;
    int64_t n = 0;
    switch (which) {
        case B2_ACT_HIDDEN: src = c->x; n = I.d_model; break;
        case B2_ACT_Q: src = c->q; n = (int64_t)I.n_head * I.head_dim; break;
        case B2_ACT_ATTN_OUT: src = c->attn_out; n = (int64_t)I.n_head * I.head_dim; break;
        case B2_ACT_FFN_H: src = c->h; n = I.d_ff; break;
        case B2_ACT_LOGITS: src = c->logits; n = I.vocab; break;
        default: set_error("b2_ctx_get_activation: unknown buffer %d", which); return B2_EINVAL;
    }
    if (cap < n) {
        set_error("b2_ctx_get_activation: buffer %d holds %lld floats, caller provided %lld", which, (long long)n, (long long)cap);
        return B2_ERANGE;
    }
    B2_CUDA_CHECK(cudaMemcpyAsync(out_host, src, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    return (int)n;
}

extern "C" int b2_ctx_get_kv(b2_ctx* c, int layer, int pos, uint16_t* k_out_host, uint16_t* v_out_host) {
    if (!c || !k_out_host || !v_out_host) return B2_EINVAL;
    const b2_model_info& I = c->m->info;
    if (layer < 0 || layer >= I.n_layer || pos < 0 || pos >= c->n_ctx) {
        set_error("b2_ctx_get_kv: layer %d / position %d out of range", layer, pos);
        return B2_ERANGE;
    }
    const size_t kv_layer = (size_t)I.n_kv * c->kv_tokens * I.head_dim;
    for (int h = 0; h < I.n_kv; h++) {
        const size_t off = layer * kv_layer + ((size_t)h * c->kv_tokens + pos) * I.head_dim;   // slot 0: identity pages
        B2_CUDA_CHECK(cudaMemcpyAsync(k_out_host + (size_t)h * I.head_dim, c->k_cache + off, I.head_dim * 2, cudaMemcpyDeviceToHost, c->stream));
        B2_CUDA_CHECK(cudaMemcpyAsync(v_out_host + (size_t)h * I.head_dim, c->v_cache + off, I.head_dim * 2, cudaMemcpyDeviceToHost, c->stream));
    }
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    return 0;
}

// ---- measurement helpers (bench.py): CUDA events on the context's own stream ----------------------
extern "C" int b2_ctx_time_steps(b2_ctx* c, int n_steps, float* ms_out) {
    if (!c || !ms_out || n_steps <= 0) return B2_EINVAL;
    B2_CUDA_CHECK(cudaSetDevice(c->m->device));
    cudaEvent_t e0, e1;
    B2_CUDA_CHECK(cudaEventCreate(&e0));
    B2_CUDA_CHECK(cudaEventCreate(&e1));
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    B2_CUDA_CHECK(cudaEventRecord(e0, c->stream));
    int rc = b2_ctx_enqueue_steps(c, n_steps);
    if (rc == 0) {
        cudaError_t e = cudaEventRecord(e1, c->stream);
        if (e == cudaSuccess) e = cudaEventSynchronize(e1);
        if (e == cudaSuccess) e = cudaEventElapsedTime(ms_out, e0, e1);
        if (e != cudaSuccess) {
            set_error("b2_ctx_time_steps: %s", cudaGetErrorString(e));
            rc = B2_ECUDA;
        }
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (rc == 0) dump_step_trace(c);
    return rc;
}

extern "C" int b2_ctx_time_matvec_only(b2_ctx* c, int reps, float* ms_out, uint64_t* bytes_per_rep_out, int* launches_per_rep_out) {
    if (!c || !ms_out || reps <= 0) return B2_EINVAL;
    B2_CUDA_CHECK(cudaSetDevice(c->m->device));
    cudaEvent_t e0, e1;
    B2_CUDA_CHECK(cudaEventCreate(&e0));
    B2_CUDA_CHECK(cudaEventCreate(&e1));
    int rc = enqueue_step(c, c->stream, false, true);  // warm
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    B2_CUDA_CHECK(cudaEventRecord(e0, c->stream));
    for (int i = 0; i < reps && rc == 0; i++) rc = enqueue_step(c, c->stream, false, true);
    if (rc == 0) {
        cudaError_t e = cudaEventRecord(e1, c->stream);
        if (e == cudaSuccess) e = cudaEventSynchronize(e1);
        if (e == cudaSuccess) e = cudaEventElapsedTime(ms_out, e0, e1);
        if (e != cudaSuccess) {
            set_error("b2_ctx_time_matvec_only: %s", cudaGetErrorString(e));
            rc = B2_ECUDA;
        }
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    // weight bytes the mat-vec launches of one step read: every matrix once (norm vectors and biases excluded)
    if (bytes_per_rep_out) {
        const b2_model_info& I = c->m->info;
        uint64_t b = c->m->output.rows * (uint64_t)row_bytes(c->m->output.type, c->m->output.cols);
        for (auto& L : c->m->layers)
            for (const b2_model::W* w : {&L.wq, &L.wk, &L.wv, &L.wo, &L.gate, &L.up, &L.down}) b += (uint64_t)w->rows * row_bytes(w->type, w->cols);
        (void)I;
        *bytes_per_rep_out = b;
    }
    if (launches_per_rep_out) *launches_per_rep_out = c->matvec_launches;
    // the residual epilogues ran on stale activations: put the context back to a clean sequence start
    B2_CUDA_CHECK(cudaMemsetAsync(c->d_pos, 0, 4, c->stream));
    B2_CUDA_CHECK(cudaStreamSynchronize(c->stream));
    c->pos = 0;
    return rc;
}
