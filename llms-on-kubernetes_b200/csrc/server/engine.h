// engine.h -- one model, one decode context with several sequence slots, and the scheduler that turns
// concurrent HTTP requests into batches (continuous batching; the reference's Services and
// gateways deliver requests concurrently, ramalama-models/helm-chart/templates/model-services.yaml:16-17).
//
// Connection threads call generate(): it queues a job and relays the text pieces the scheduler produces.
// The scheduler thread owns the GPU context.  Every iteration it builds one batch of up to max_cols columns (32, or
// 256 with the model's fp16 copy resident) for
// b2_ctx_decode_batch: one column per sequence that is generating, the rest filled with prompt tokens of
// the sequences that are still being prefilled.  A lone greedy request in slot 0 takes the device-resident
// single-sequence loop instead (a few tokens per iteration, so a newcomer waits a few milliseconds at most).
// Each slot keeps the token ids whose K/V rows it holds: a new request reuses the longest common prefix.
#pragma once
#include <algorithm>
#include <atomic>
#include <cmath>
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "../../../include/b2.h"
#include "tokenizer.h"

namespace b2srv {

struct GenParams {
    int max_tokens = 256;
    float temperature = 0.8f, top_p = 0.95f, min_p = 0.05f;   // llama-server's defaults when the request omits them
    int top_k = 40;
    uint64_t seed = 0xFFFFFFFFull;
    std::vector<std::string> stop;
    bool ignore_eos = false;
};
struct GenResult {
    std::vector<int32_t> tokens;
    std::string text, finish = "length";
    int prompt_tokens = 0;
    std::string error;
    int error_status = 0;
};

struct Engine {
    b2_model* model = nullptr;
    b2_ctx* ctx = nullptr;
    b2_model_info info{};
    Tokenizer tok;
    std::string alias, model_path;
    std::atomic<bool> ready{false};
    std::string load_error;

    // ---- one request in flight ----
    struct Job {
        const std::vector<int32_t>* prompt = nullptr;
        GenParams p;
        // scheduler-owned
        int slot = -1;
        size_t fed = 0;            // prompt tokens whose K/V rows are in the slot
        int32_t next = 0;          // sampled, not yet fed to the model
        bool have_next = false, greedy = true, finished = false;
        int max_new = 0;
        std::mt19937_64 rng;
        uint64_t seed_used = 0;
        std::string pending;       // text held back: it may be the start of a stop string or an incomplete UTF-8 sequence
        size_t max_stop = 0;
        GenResult res;
        // hand-off to the connection thread
        std::mutex m;
        std::condition_variable cv;
        std::deque<std::string> pieces;
        bool done = false;
        std::atomic<bool> cancel{false};
    };

    int n_slots = 1;
    std::vector<std::vector<int32_t>> slot_cached;   // per slot: tokens at positions [0, size) of its KV cache
    std::vector<Job*> slot_job;
    std::vector<char> slot_samples;   // per slot: a device sampler is set (b2_ctx_set_sampler)
    bool dev_sampler = true;
    std::mutex qmu;
    std::condition_variable qcv;
    std::deque<Job*> waiting;
    std::thread sched;
    std::vector<float> logits;
    bool batch_ok = true;   // false: the model's types are not on the batched path -> one request at a time
    // counters behind GET /metrics (Prometheus text; llama.cpp's metric names where it has one)
    std::atomic<uint64_t> m_requests{0}, m_prompt_tokens{0}, m_prompt_reused{0}, m_predicted{0}, m_passes{0}, m_pass_cols{0};
    std::atomic<int> m_processing{0}, m_deferred{0};
    std::atomic<uint64_t> m_host_sampled{0}, m_dev_sampler_requests{0};   // tokens drawn on the host from copied logits; requests whose slot draws on the device
    int max_cols = 32;      // widest batch per pass (b2_ctx_max_batch: 32 on the int8 kernel, 256 on the tcgen05 pass)
    bool tc_prefill = false;   // long prompts go through b2_ctx_prefill (whole chunks on the tensor-core prompt path)

    int n_ctx() const { return b2_ctx_n_ctx(ctx); }

    bool load(const std::string& path, int n_ctx_req, int slots_req) {
        int rc = b2_model_load(path.c_str(), 0, 0, 1, &model);
        if (rc) { load_error = b2_last_error(); return false; }
        b2_model_get_info(model, &info);
        std::string err;
        if (!tok.load(b2_model_gguf(model), &err)) { load_error = err; return false; }
        n_slots = std::max(1, std::min(slots_req, 256));
        rc = b2_ctx_create_slots(model, n_ctx_req, n_slots, &ctx);
        if (rc) { load_error = b2_last_error(); return false; }
        // probe the batched path once: models with other matrix types keep the single-sequence path
        {
            int32_t t = 0, p0 = 0, out = 0;
            if (b2_ctx_decode_batch(ctx, &t, nullptr, &p0, 1, 1, &out, nullptr) != 0) {
                batch_ok = false;
                n_slots = 1;
            }
        }
        if (batch_ok) {
            const int mb = b2_ctx_max_batch(ctx);   // builds the resident fp16 copy when it fits
            if (mb > max_cols) {
                max_cols = mb;
                tc_prefill = !getenv("B2_NO_TC_PREFILL");
            }
        }
        slot_cached.assign(n_slots, {});
        slot_job.assign(n_slots, nullptr);
        slot_samples.assign(n_slots, 0);
        dev_sampler = !getenv("B2_HOST_SAMPLER");   // B2_HOST_SAMPLER=1: logits to the host, Engine::sample (the reference for the tests)
        logits.resize(info.vocab);
        sched = std::thread([this] { run_scheduler(); });
        sched.detach();
        ready = true;
        return true;
    }

    int32_t sample(const GenParams& p, std::mt19937_64& rng) {
        m_host_sampled++;
        const int V = info.vocab;
        std::vector<std::pair<float, int32_t>> c(V);
        for (int i = 0; i < V; i++) c[i] = {logits[i], i};
        int k = p.top_k > 0 && p.top_k < V ? p.top_k : V;
        std::partial_sort(c.begin(), c.begin() + k, c.end(), [](auto& a, auto& b) { return a.first > b.first || (a.first == b.first && a.second < b.second); });
        c.resize(k);
        const float mx = c[0].first;
        double sum = 0;
        std::vector<double> pr(k);
        for (int i = 0; i < k; i++) { pr[i] = std::exp((double)(c[i].first - mx) / p.temperature); sum += pr[i]; }
        for (auto& x : pr) x /= sum;
        int keep = k;
        if (p.min_p > 0) {
            keep = 0;
            while (keep < k && pr[keep] >= p.min_p * pr[0]) keep++;
            keep = std::max(keep, 1);
        }
        if (p.top_p < 1.0f) {
            double acc = 0;
            int n = 0;
            while (n < keep) { acc += pr[n++]; if (acc >= p.top_p) break; }
            keep = std::max(n, 1);
        }
        double tot = 0;
        for (int i = 0; i < keep; i++) tot += pr[i];
        double r = std::uniform_real_distribution<double>(0.0, tot)(rng);
        for (int i = 0; i < keep; i++) { r -= pr[i]; if (r <= 0) return c[i].second; }
        return c[keep - 1].second;
    }

    // on_text(piece) is called (on the caller's thread) as text becomes final; returning false aborts (client gone)
    GenResult generate(const std::vector<int32_t>& prompt, const GenParams& p, const std::function<bool(const std::string&)>& on_text) {
        Job job;
        job.prompt = &prompt;
        job.p = p;
        job.res.prompt_tokens = (int)prompt.size();
        const int nctx = n_ctx();
        if (prompt.empty()) { job.res.error = "the prompt is empty"; job.res.error_status = 400; return job.res; }
        if ((int)prompt.size() >= nctx) {
            job.res.error = "the prompt (" + std::to_string(prompt.size()) + " tokens) does not fit the context size " + std::to_string(nctx);
            job.res.error_status = 400;
            return job.res;
        }
        for (int32_t t : prompt)
            if (t < 0 || t >= info.vocab) { job.res.error = "token id out of range in prompt"; job.res.error_status = 400; return job.res; }
        job.max_new = std::min(p.max_tokens < 0 ? nctx : p.max_tokens, nctx - (int)prompt.size());
        job.greedy = p.temperature <= 0.0f || p.top_k == 1;
        job.seed_used = p.seed == 0xFFFFFFFFull ? ((uint64_t)std::random_device{}() << 32 | std::random_device{}()) : p.seed;
        job.rng.seed(job.seed_used);
        for (auto& s : p.stop) job.max_stop = std::max(job.max_stop, s.size());
        m_requests++;
        m_deferred++;
        {
            std::lock_guard<std::mutex> lk(qmu);
            waiting.push_back(&job);
        }
        qcv.notify_one();
        // relay text until the scheduler is done with the job
        std::unique_lock<std::mutex> lk(job.m);
        for (;;) {
            job.cv.wait(lk, [&] { return job.done || !job.pieces.empty(); });
            while (!job.pieces.empty()) {
                std::string piece = std::move(job.pieces.front());
                job.pieces.pop_front();
                lk.unlock();
                if (on_text && !job.cancel && !on_text(piece)) job.cancel = true;
                lk.lock();
            }
            if (job.done) break;
        }
        return job.res;
    }

    // ---- scheduler thread ------------------------------------------------------------------------------
    void push_piece(Job& j, const std::string& piece) {
        j.res.text += piece;
        {
            std::lock_guard<std::mutex> lk(j.m);
            j.pieces.push_back(piece);
        }
        j.cv.notify_one();
    }
    // emit everything of j.pending except a tail that could still grow into a stop string or a UTF-8 sequence
    void flush(Job& j, bool final) {
        std::string& pending = j.pending;
        const size_t hold = final ? 0 : (j.max_stop ? j.max_stop - 1 : 0);
        size_t emit = pending.size() > hold ? pending.size() - hold : 0;
        while (!final && emit > 0 && emit < pending.size() && ((unsigned char)pending[emit] & 0xC0) == 0x80) emit--;
        if (!final && emit == pending.size()) {
            // ... nor a lead byte whose continuation bytes have not been generated yet
            for (size_t back = 1; back <= 3 && back <= emit; back++) {
                const unsigned char c = (unsigned char)pending[emit - back];
                if ((c & 0xC0) == 0x80) continue;
                const size_t need = c >= 0xF0 ? 4 : c >= 0xE0 ? 3 : c >= 0xC0 ? 2 : 1;
                if (need > back) emit -= back;
                break;
            }
        }
        if (!emit) return;
        const std::string piece = pending.substr(0, emit);
        pending.erase(0, emit);
        push_piece(j, piece);
    }
    void finish(Job& j, const char* why) {
        if (j.finished) return;
        if (why) j.res.finish = why;
        flush(j, true);
        j.finished = true;
        j.have_next = false;
    }
    // j.next has been sampled but not yet fed: account for it; returns false when the job is finished
    bool accept_next(Job& j) {
        if (j.cancel) { finish(j, "abort"); return false; }
        if (!j.p.ignore_eos && tok.is_eog(j.next)) { finish(j, "stop"); return false; }
        if ((int)j.res.tokens.size() >= j.max_new) { finish(j, nullptr); return false; }
        j.res.tokens.push_back(j.next);
        m_predicted++;
        tok.decode_append(j.next, j.pending);
        for (auto& s : j.p.stop) {
            const size_t at = s.empty() ? std::string::npos : j.pending.find(s);
            if (at != std::string::npos) {
                j.pending.erase(at);
                finish(j, "stop");
                return false;
            }
        }
        flush(j, false);
        if ((int)j.res.tokens.size() >= j.max_new) { finish(j, nullptr); return false; }
        if ((int)slot_cached[j.slot].size() + 1 >= n_ctx()) { finish(j, nullptr); return false; }
        j.have_next = true;
        return true;
    }
    void fail(Job& j, const std::string& msg) {
        j.res.error = msg;
        j.res.error_status = 500;
        slot_cached[j.slot].clear();
        j.finished = true;
    }
    void retire_finished() {
        for (int s = 0; s < n_slots; s++) {
            Job* j = slot_job[s];
            if (!j || !j->finished) continue;
            slot_job[s] = nullptr;   // the slot keeps its tokens: the next request may share the prefix
            m_processing--;
            // notify under the lock: the waiter cannot return (and destroy the job) before we are done with it
            std::lock_guard<std::mutex> lk(j->m);
            j->done = true;
            j->cv.notify_one();
        }
    }
    void admit() {
        std::lock_guard<std::mutex> lk(qmu);
        while (!waiting.empty()) {
            Job* j = waiting.front();
            // the free slot sharing the longest prefix with the prompt (slot 0 on ties: it has the fast single-sequence path)
            int best = -1;
            size_t best_cp = 0;
            for (int s = 0; s < n_slots; s++) {
                if (slot_job[s]) continue;
                size_t cp = 0;
                const auto& c = slot_cached[s];
                while (cp < c.size() && cp < j->prompt->size() && c[cp] == (*j->prompt)[cp]) cp++;
                if (best < 0 || cp > best_cp) { best = s; best_cp = cp; }
            }
            if (best < 0) break;   // all slots busy: the job waits
            waiting.pop_front();
            if (best_cp >= j->prompt->size()) best_cp = j->prompt->size() - 1;   // always re-evaluate the last prompt token
            j->slot = best;
            j->fed = best_cp;
            slot_cached[best].resize(best_cp);
            b2_ctx_slot_truncate(ctx, best, (int)best_cp);   // paged KV: the pages past the shared prefix go back to the pool
            slot_job[best] = j;
            // sampling happens on the device (b2_ctx_set_sampler): the slot draws with this request's parameters, and from
            // here on the request is handled like a greedy one -- every call returns its next token, no logits come back
            if (dev_sampler) {
                if (!j->greedy) {
                    const b2_sampling sp{j->p.temperature, j->p.top_p > 0.0f && j->p.top_p <= 1.0f ? j->p.top_p : 1.0f,
                                         j->p.min_p < 0.0f ? 0.0f : (j->p.min_p > 1.0f ? 1.0f : j->p.min_p), j->p.top_k, j->seed_used};
                    if (b2_ctx_set_sampler(ctx, best, &sp) == 0) { j->greedy = true; slot_samples[best] = 1; m_dev_sampler_requests++; }
                } else if (slot_samples[best]) {
                    b2_ctx_set_sampler(ctx, best, nullptr);
                    slot_samples[best] = 0;
                }
            }
            m_deferred--;
            m_processing++;
            m_prompt_reused += best_cp;
            m_prompt_tokens += j->prompt->size() - best_cp;
        }
    }

    // a lone greedy sequence in slot 0: the persistent per-token kernel, a few tokens per call
    void single_greedy_chunk(Job& j) {
        auto& cached = slot_cached[0];
        int32_t t = 0;
        int rc = b2_ctx_decode(ctx, &j.next, 1, (int)cached.size(), nullptr, &t);
        if (rc) { fail(j, b2_last_error()); return; }
        cached.push_back(j.next);
        j.next = t;
        j.have_next = false;
        if (!accept_next(j)) return;
        const int k = std::min({7, j.max_new - (int)j.res.tokens.size(), n_ctx() - (int)cached.size() - 1});
        if (k <= 0) return;
        int32_t chunk[8];
        rc = b2_ctx_generate_greedy(ctx, k, chunk);   // feeds j.next, chunk[0..k-2]; chunk[k-1] is sampled, not fed
        if (rc) { fail(j, b2_last_error()); return; }
        cached.push_back(j.next);
        for (int i = 0; i < k; i++) {
            j.next = chunk[i];
            j.have_next = false;
            if (!accept_next(j)) return;   // rows past this token in the cache are simply not recorded
            if (i < k - 1) cached.push_back(chunk[i]);
        }
    }

    // the KV pool is out of pages: idle slots give up the prefixes they were keeping for reuse
    bool reclaim_idle_slots() {
        bool any = false;
        for (int s = 1; s < n_slots; s++)
            if (!slot_job[s] && !slot_cached[s].empty()) {
                slot_cached[s].clear();
                b2_ctx_slot_truncate(ctx, s, 0);
                any = true;
            }
        return any;
    }

    void run_scheduler() {
        std::vector<int32_t> b_tok, b_slot, b_pos, b_out;
        std::vector<Job*> b_job;
        std::vector<char> b_last;   // column whose logits pick the job's next token
        for (;;) {
            {
                std::unique_lock<std::mutex> lk(qmu);
                qcv.wait(lk, [&] {
                    if (!waiting.empty()) return true;
                    for (Job* j : slot_job)
                        if (j) return true;
                    return false;
                });
            }
            admit();
            int n_active = 0;
            Job* only = nullptr;
            for (Job* j : slot_job)
                if (j) { n_active++; only = j; }
            if (!n_active) continue;
            for (Job* j : slot_job)
                if (j && j->cancel && !j->finished) finish(*j, "abort");
            bool queue_empty;
            {
                std::lock_guard<std::mutex> lk(qmu);
                queue_empty = waiting.empty();
            }
            if (n_active == 1 && queue_empty && only->slot == 0 && only->greedy && only->have_next && !only->finished) {
                single_greedy_chunk(*only);
                retire_finished();
                continue;
            }
            if (!batch_ok) {
                // single-sequence model: the whole prompt through b2_ctx_decode, then token by token
                Job& j = *only;
                if (!j.finished && !j.have_next) {
                    int32_t t = 0;
                    const int rc = b2_ctx_decode(ctx, j.prompt->data() + j.fed, (int)(j.prompt->size() - j.fed), (int)j.fed,
                                                 j.greedy ? nullptr : logits.data(), &t);
                    if (rc) fail(j, b2_last_error());
                    else {
                        slot_cached[0] = *j.prompt;
                        j.fed = j.prompt->size();
                        j.next = j.greedy ? t : sample(j.p, j.rng);
                        accept_next(j);
                    }
                } else if (!j.finished) {
                    int32_t t = 0;
                    const int rc = b2_ctx_decode(ctx, &j.next, 1, (int)slot_cached[0].size(), j.greedy ? nullptr : logits.data(), &t);
                    if (rc) fail(j, b2_last_error());
                    else {
                        slot_cached[0].push_back(j.next);
                        j.next = j.greedy ? t : sample(j.p, j.rng);
                        j.have_next = false;
                        accept_next(j);
                    }
                }
                retire_finished();
                continue;
            }
            // ---- long prompts: the remainders of as many waiting prompts as fit one pass of the tensor-core prompt path
            // (b2_ctx_prefill_multi: up to 2048 rows); a prompt longer than that goes alone, in chunks (b2_ctx_prefill) ----
            if (tc_prefill) {
                std::vector<Job*> todo;
                for (Job* j : slot_job)
                    if (j && !j->finished && !j->have_next && j->prompt->size() - j->fed >= 64) todo.push_back(j);
                const size_t max_rows = (size_t)std::min(2048, n_ctx());
                size_t at = 0;
                while (at < todo.size()) {
                    std::vector<Job*> grp;
                    std::vector<int32_t> g_slot, g_pos0, g_len, g_tok, g_out;
                    size_t rows = 0;
                    while (at < todo.size() && grp.size() < 64) {
                        Job* j = todo[at];
                        const size_t rem = j->prompt->size() - j->fed;
                        if (!grp.empty() && rows + rem > max_rows) break;
                        grp.push_back(j);
                        at++;
                        rows += rem;
                        if (rem > max_rows) break;   // alone: chunked inside b2_ctx_prefill
                        g_slot.push_back(j->slot);
                        g_pos0.push_back((int32_t)j->fed);
                        g_len.push_back((int32_t)rem);
                        g_tok.insert(g_tok.end(), j->prompt->begin() + j->fed, j->prompt->end());
                    }
                    const bool single = grp.size() == 1;   // (also every sampled request: its logits come back alone)
                    g_out.assign(grp.size(), 0);
                    auto call = [&]() -> int {
                        if (single) {
                            Job* j = grp[0];
                            return b2_ctx_prefill(ctx, j->slot, j->prompt->data() + j->fed, (int)(j->prompt->size() - j->fed), (int)j->fed, &g_out[0],
                                                  j->greedy ? nullptr : logits.data());
                        }
                        return b2_ctx_prefill_multi(ctx, (int)grp.size(), g_slot.data(), g_pos0.data(), g_len.data(), g_tok.data(), g_out.data(), nullptr);
                    };
                    int rc = call();
                    if (rc == B2_ENOMEM && reclaim_idle_slots()) rc = call();
                    if (rc) {
                        const std::string msg = b2_last_error();
                        for (Job* j : grp) fail(*j, msg);
                        continue;
                    }
                    for (size_t i = 0; i < grp.size(); i++) {
                        Job* j = grp[i];
                        auto& cached = slot_cached[j->slot];
                        cached.insert(cached.end(), j->prompt->begin() + j->fed, j->prompt->end());
                        j->fed = j->prompt->size();
                        if (j->greedy) j->next = g_out[i];
                        else if (single) j->next = sample(j->p, j->rng);
                        else {
                            // a sampled request inside a group: its last prompt token is fed again through the batch below, whose
                            // per-column logits reach the host sampler
                            cached.pop_back();
                            j->fed--;
                            continue;
                        }
                        j->have_next = false;
                        accept_next(*j);
                    }
                }
            }
            // ---- one batch: a column per generating sequence, the rest prompt tokens ----
            b_tok.clear(); b_slot.clear(); b_pos.clear(); b_job.clear(); b_last.clear();
            const int cap = max_cols;
            for (Job* j : slot_job)
                if (j && !j->finished && j->have_next && (int)b_tok.size() < cap) {
                    b_tok.push_back(j->next);
                    b_slot.push_back(j->slot);
                    b_pos.push_back((int32_t)slot_cached[j->slot].size());
                    b_job.push_back(j);
                    b_last.push_back(1);
                }
            for (Job* j : slot_job) {
                if (!j || j->finished || j->have_next || j->fed >= j->prompt->size()) continue;
                const int room = cap - (int)b_tok.size();
                if (room <= 0) break;
                const int n = (int)std::min<size_t>((size_t)room, j->prompt->size() - j->fed);
                for (int i = 0; i < n; i++) {
                    b_tok.push_back((*j->prompt)[j->fed + i]);
                    b_slot.push_back(j->slot);
                    b_pos.push_back((int32_t)(j->fed + i));
                    b_job.push_back(j);
                    b_last.push_back(j->fed + i + 1 == j->prompt->size());
                }
            }
            const int n = (int)b_tok.size();
            if (n == 0) { retire_finished(); continue; }
            b_out.resize(n);
            m_passes++;
            m_pass_cols += (uint64_t)n;
            int rc = b2_ctx_decode_batch(ctx, b_tok.data(), b_slot.data(), b_pos.data(), n, 0, b_out.data(), nullptr);
            if (rc == B2_ENOMEM && reclaim_idle_slots())
                rc = b2_ctx_decode_batch(ctx, b_tok.data(), b_slot.data(), b_pos.data(), n, 0, b_out.data(), nullptr);
            if (rc) {
                const std::string msg = b2_last_error();
                for (Job* j : b_job)
                    if (!j->finished) fail(*j, msg);
                retire_finished();
                continue;
            }
            for (int c = 0; c < n; c++) {
                Job& j = *b_job[c];
                if (j.finished) continue;
                slot_cached[j.slot].push_back(b_tok[c]);
                if ((size_t)b_pos[c] >= j.fed && j.fed < j.prompt->size()) j.fed = (size_t)b_pos[c] + 1;
                if (!b_last[c]) continue;
                if (j.greedy) j.next = b_out[c];
                else {
                    if (b2_ctx_batch_logits(ctx, c, logits.data()) != 0) { fail(j, b2_last_error()); continue; }
                    j.next = sample(j.p, j.rng);
                }
                j.have_next = false;
                accept_next(j);
            }
            retire_finished();
        }
    }
};

}  // namespace b2srv
