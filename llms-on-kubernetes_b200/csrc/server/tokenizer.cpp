// tokenizer.cpp -- SentencePiece-style ("llama") and byte-level BPE ("gpt2") tokenisation driven by the
// GGUF's tokenizer.ggml.* metadata, plus the chat-template families the reference's example models use
// (Zephyr for TinyLlama, ChatML for Qwen, Llama-3, Llama-2/Mistral [INST]).  Replaces the tokenizer stage
// of the llama-server the chart launches.
//
// Fidelity: the SPM path follows llama.cpp's algorithm (UTF-8 symbols, highest-score bigram merged first,
// <0xXX> byte fallback).  The BPE path applies the GGUF's merge ranks over a GPT-2-style pre-split that is an
// approximation of the per-model regexes (exact pre-tokenisers are "next").  Templates are
// recognised by their markers, not interpreted as Jinja.
#include "tokenizer.h"

#include <algorithm>
#include <cstring>
#include <queue>

namespace b2srv {

static bool kv_string(const b2_gguf* g, const char* key, std::string* out) {
    const int64_t i = b2_gguf_find_kv(g, key);
    const char* s;
    uint64_t n;
    if (i < 0 || b2_gguf_kv_str(g, i, &s, &n) != 0) return false;
    out->assign(s, n);
    return true;
}
static bool kv_int(const b2_gguf* g, const char* key, int64_t* out) {
    const int64_t i = b2_gguf_find_kv(g, key);
    return i >= 0 && b2_gguf_kv_i64(g, i, out) == 0;
}

bool Tokenizer::load(const b2_gguf* g, std::string* err) {
    if (!kv_string(g, "tokenizer.ggml.model", &model)) model = "llama";
    const int64_t ti = b2_gguf_find_kv(g, "tokenizer.ggml.tokens");
    if (ti < 0 || b2_gguf_kv_type(g, ti) != B2_KV_ARRAY) {
        if (err) *err = "GGUF has no tokenizer.ggml.tokens";
        return false;
    }
    const int64_t n = b2_gguf_kv_array_len(g, ti);
    tokens.resize(n);
    for (int64_t j = 0; j < n; j++) {
        const char* s;
        uint64_t len;
        if (b2_gguf_kv_arr_str(g, ti, j, &s, &len) != 0) {
            if (err) *err = "malformed tokenizer.ggml.tokens";
            return false;
        }
        tokens[j].assign(s, len);
    }
    scores.assign(n, 0.0f);
    types.assign(n, 1);
    const int64_t si = b2_gguf_find_kv(g, "tokenizer.ggml.scores");
    if (si >= 0 && b2_gguf_kv_array_len(g, si) == n)
        for (int64_t j = 0; j < n; j++) {
            double d;
            if (b2_gguf_kv_arr_f64(g, si, j, &d) == 0) scores[j] = (float)d;
        }
    const int64_t yi = b2_gguf_find_kv(g, "tokenizer.ggml.token_type");
    if (yi >= 0 && b2_gguf_kv_array_len(g, yi) == n)
        for (int64_t j = 0; j < n; j++) {
            int64_t v;
            if (b2_gguf_kv_arr_i64(g, yi, j, &v) == 0) types[j] = (int32_t)v;
        }
    tok2id.reserve(n * 2);
    for (int32_t j = (int32_t)n - 1; j >= 0; j--) tok2id[tokens[j]] = j;   // lowest id wins duplicates
    const int64_t mi = b2_gguf_find_kv(g, "tokenizer.ggml.merges");
    if (mi >= 0) {
        const int64_t nm = b2_gguf_kv_array_len(g, mi);
        merge_rank.reserve(nm * 2);
        for (int64_t j = 0; j < nm; j++) {
            const char* s;
            uint64_t len;
            if (b2_gguf_kv_arr_str(g, mi, j, &s, &len) == 0) merge_rank.emplace(std::string(s, len), (int32_t)j);
        }
    }
    int64_t v;
    if (kv_int(g, "tokenizer.ggml.bos_token_id", &v)) bos_id = (int32_t)v;
    if (kv_int(g, "tokenizer.ggml.eos_token_id", &v)) eos_id = (int32_t)v;
    if (kv_int(g, "tokenizer.ggml.eot_token_id", &v)) eot_id = (int32_t)v;
    if (kv_int(g, "tokenizer.ggml.unknown_token_id", &v)) unk_id = (int32_t)v;
    if (kv_int(g, "tokenizer.ggml.add_bos_token", &v)) add_bos = v != 0;
    else add_bos = model == "llama";
    if (kv_int(g, "tokenizer.ggml.add_space_prefix", &v)) add_space_prefix = v != 0;
    kv_string(g, "tokenizer.chat_template", &tmpl);

    for (int b = 0; b < 256; b++) {
        char name[8];
        snprintf(name, sizeof name, "<0x%02X>", b);
        auto it = tok2id.find(name);
        byte_tok[b] = it == tok2id.end() ? -1 : it->second;
    }
    // control (3) and user-defined (4) tokens are matched verbatim in chat prompts
    for (int32_t j = 0; j < (int32_t)n; j++)
        if ((types[j] == 3 || types[j] == 4) && !tokens[j].empty()) specials.emplace_back(tokens[j], j);
    std::sort(specials.begin(), specials.end(), [](auto& a, auto& b) { return a.first.size() > b.first.size(); });
    for (const char* name : {"<|eot_id|>", "<|im_end|>", "<|end|>", "<|endoftext|>", "<|end_of_text|>", "</s>"}) {
        auto it = tok2id.find(name);
        if (it != tok2id.end() && types[it->second] != 1) eog_extra[it->second] = true;
    }
    // template family by its markers
    if (tmpl.find("<|im_start|>") != std::string::npos) tmpl_kind = "chatml";
    else if (tmpl.find("<|start_header_id|>") != std::string::npos) tmpl_kind = "llama3";
    else if (tmpl.find("<|user|>") != std::string::npos) tmpl_kind = tmpl.find("<|end|>") != std::string::npos ? "phi3" : "zephyr";
    else if (tmpl.find("[INST]") != std::string::npos) tmpl_kind = "llama2";
    else tmpl_kind = tok2id.count("<|im_start|>") ? "chatml" : "plain";
    return true;
}

// ---- SPM -----------------------------------------------------------------------------------------
namespace {
struct Sym {
    int prev, next;
    size_t off, len;
};
struct Bigram {
    int left, right;
    float score;
    size_t len;
    bool operator<(const Bigram& o) const { return score < o.score || (score == o.score && left > o.left); }
};
size_t utf8_len(unsigned char c) { return c < 0x80 ? 1 : c < 0xE0 ? 2 : c < 0xF0 ? 3 : 4; }
}  // namespace

void Tokenizer::encode_spm(const std::string& raw, std::vector<int32_t>& out) const {
    // spaces become U+2581; a leading one is added when the model asks for it
    std::string text;
    if (add_space_prefix && !raw.empty()) text = "\xE2\x96\x81";
    for (char c : raw) {
        if (c == ' ') text += "\xE2\x96\x81";
        else text.push_back(c);
    }
    std::vector<Sym> syms;
    for (size_t off = 0; off < text.size();) {
        size_t len = std::min(utf8_len((unsigned char)text[off]), text.size() - off);
        syms.push_back({(int)syms.size() - 1, (int)syms.size() + 1, off, len});
        off += len;
    }
    if (syms.empty()) return;
    syms.back().next = -1;
    std::priority_queue<Bigram> work;
    auto try_add = [&](int l, int r) {
        if (l < 0 || r < 0) return;
        const std::string piece = text.substr(syms[l].off, syms[l].len + syms[r].len);
        auto it = tok2id.find(piece);
        if (it == tok2id.end()) return;
        work.push({l, r, scores[it->second], piece.size()});
    };
    for (int i = 1; i < (int)syms.size(); i++) try_add(i - 1, i);
    while (!work.empty()) {
        const Bigram bg = work.top();
        work.pop();
        Sym& L = syms[bg.left];
        Sym& R = syms[bg.right];
        if (L.len == 0 || R.len == 0 || L.len + R.len != bg.len) continue;   // stale entry
        L.len += R.len;
        R.len = 0;
        L.next = R.next;
        if (R.next >= 0) syms[R.next].prev = bg.left;
        try_add(L.prev, bg.left);
        try_add(bg.left, L.next);
    }
    for (int i = 0; i >= 0; i = syms[i].next) {
        if (syms[i].len == 0) continue;
        const std::string piece = text.substr(syms[i].off, syms[i].len);
        auto it = tok2id.find(piece);
        if (it != tok2id.end()) {
            out.push_back(it->second);
        } else {
            for (unsigned char c : piece) {
                if (byte_tok[c] >= 0) out.push_back(byte_tok[c]);
                else if (unk_id >= 0) out.push_back(unk_id);
            }
        }
    }
}

// ---- byte-level BPE --------------------------------------------------------------------------------
namespace {
// GPT-2's bytes <-> unicode table
const std::vector<std::string>& byte_to_unicode() {
    static std::vector<std::string> tab = [] {
        std::vector<std::string> t(256);
        std::vector<int> cps(256, -1);
        int n = 0;
        for (int b = 0; b < 256; b++) {
            if ((b >= 33 && b <= 126) || (b >= 161 && b <= 172) || (b >= 174)) cps[b] = b;
        }
        for (int b = 0; b < 256; b++)
            if (cps[b] < 0) cps[b] = 256 + n++;
        for (int b = 0; b < 256; b++) {
            const int cp = cps[b];
            if (cp < 0x80) t[b] = std::string(1, (char)cp);
            else { t[b].push_back((char)(0xC0 | (cp >> 6))); t[b].push_back((char)(0x80 | (cp & 0x3F))); }
        }
        return t;
    }();
    return tab;
}
const std::unordered_map<std::string, unsigned char>& unicode_to_byte() {
    static std::unordered_map<std::string, unsigned char> m = [] {
        std::unordered_map<std::string, unsigned char> r;
        const auto& t = byte_to_unicode();
        for (int b = 0; b < 256; b++) r[t[b]] = (unsigned char)b;
        return r;
    }();
    return m;
}
bool is_letter(unsigned char c) { return (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || c >= 0x80; }
bool is_digit(unsigned char c) { return c >= '0' && c <= '9'; }
bool is_space(unsigned char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }
}  // namespace

void Tokenizer::bpe_word(const std::string& word, std::vector<int32_t>& out) const {
    const auto& b2u = byte_to_unicode();
    std::vector<std::string> parts;
    for (unsigned char c : word) parts.push_back(b2u[c]);
    while (parts.size() > 1) {
        int best = -1, best_rank = INT32_MAX;
        for (size_t i = 0; i + 1 < parts.size(); i++) {
            auto it = merge_rank.find(parts[i] + " " + parts[i + 1]);
            if (it != merge_rank.end() && it->second < best_rank) { best_rank = it->second; best = (int)i; }
        }
        if (best < 0) break;
        parts[best] += parts[best + 1];
        parts.erase(parts.begin() + best + 1);
    }
    for (auto& p : parts) {
        auto it = tok2id.find(p);
        if (it != tok2id.end()) { out.push_back(it->second); continue; }
        // unknown merged piece: fall back to its single-character tokens
        for (size_t off = 0; off < p.size();) {
            const size_t len = std::min(utf8_len((unsigned char)p[off]), p.size() - off);
            auto jt = tok2id.find(p.substr(off, len));
            if (jt != tok2id.end()) out.push_back(jt->second);
            else if (unk_id >= 0) out.push_back(unk_id);
            off += len;
        }
    }
}

void Tokenizer::encode_bpe(const std::string& text, std::vector<int32_t>& out) const {
    // GPT-2 style pre-split: optional leading space + letters | digits (up to 3) | other symbols; runs of whitespace
    size_t i = 0;
    const size_t n = text.size();
    while (i < n) {
        size_t j = i;
        const unsigned char c = text[i];
        if (c == '\'' && i + 1 < n && strchr("sStTmMdD", text[i + 1])) j = i + 2;
        else if (c == '\'' && i + 2 < n && (!strncasecmp(&text[i + 1], "re", 2) || !strncasecmp(&text[i + 1], "ve", 2) || !strncasecmp(&text[i + 1], "ll", 2))) j = i + 3;
        else {
            size_t k = i;
            if (c == ' ' && k + 1 < n && !is_space(text[k + 1])) k++;
            const unsigned char d = text[k];
            if (is_letter(d)) { j = k; while (j < n && is_letter(text[j])) j++; }
            else if (is_digit(d)) { j = k; int cnt = 0; while (j < n && is_digit(text[j]) && cnt < 3) { j++; cnt++; } }
            else if (!is_space(d)) { j = k; while (j < n && !is_space(text[j]) && !is_letter(text[j]) && !is_digit(text[j])) j++; }
            else {
                j = i;
                while (j < n && is_space(text[j])) j++;
                // leave the last space for the following word
                if (j < n && j - i > 1 && text[j - 1] == ' ') j--;
            }
        }
        if (j <= i) j = i + 1;
        bpe_word(text.substr(i, j - i), out);
        i = j;
    }
}

std::vector<int32_t> Tokenizer::encode(const std::string& text, bool with_bos, bool parse_special) const {
    std::vector<int32_t> out;
    if (with_bos && add_bos && bos_id >= 0) out.push_back(bos_id);
    auto plain = [&](const std::string& s) {
        if (s.empty()) return;
        if (model == "gpt2") encode_bpe(s, out);
        else encode_spm(s, out);
    };
    if (!parse_special || specials.empty()) {
        plain(text);
        return out;
    }
    size_t start = 0, i = 0;
    while (i < text.size()) {
        bool hit = false;
        if (text[i] == '<' || text[i] == '[') {
            for (auto& sp : specials) {
                if (text.compare(i, sp.first.size(), sp.first) == 0) {
                    plain(text.substr(start, i - start));
                    out.push_back(sp.second);
                    i += sp.first.size();
                    start = i;
                    hit = true;
                    break;
                }
            }
        }
        if (!hit) i++;
    }
    plain(text.substr(start));
    return out;
}

void Tokenizer::decode_append(int32_t id, std::string& out) const {
    if (id < 0 || id >= (int32_t)tokens.size()) return;
    const int t = types[id];
    if (t == 3) return;   // control tokens are not text
    const std::string& s = tokens[id];
    if (t == 6 && s.size() == 6 && s[0] == '<' && s[1] == '0' && s[2] == 'x') {
        out.push_back((char)strtol(s.substr(3, 2).c_str(), nullptr, 16));
        return;
    }
    if (model == "gpt2") {
        const auto& u2b = unicode_to_byte();
        for (size_t off = 0; off < s.size();) {
            const size_t len = std::min(utf8_len((unsigned char)s[off]), s.size() - off);
            auto it = u2b.find(s.substr(off, len));
            if (it != u2b.end()) out.push_back((char)it->second);
            else out.append(s, off, len);
            off += len;
        }
        return;
    }
    for (size_t off = 0; off < s.size();) {
        if (s.compare(off, 3, "\xE2\x96\x81") == 0) { out.push_back(' '); off += 3; }
        else out.push_back(s[off++]);
    }
}

std::string Tokenizer::decode(const std::vector<int32_t>& ids) const {
    std::string out;
    for (int32_t id : ids) decode_append(id, out);
    return out;
}

std::string Tokenizer::apply_chat_template(const std::vector<ChatMessage>& msgs, bool gen) const {
    std::string p;
    const std::string eos = eos_id >= 0 && eos_id < (int32_t)tokens.size() ? tokens[eos_id] : "</s>";
    if (tmpl_kind == "chatml") {
        for (auto& m : msgs) p += "<|im_start|>" + m.role + "\n" + m.content + "<|im_end|>\n";
        if (gen) p += "<|im_start|>assistant\n";
    } else if (tmpl_kind == "llama3") {
        for (auto& m : msgs) p += "<|start_header_id|>" + m.role + "<|end_header_id|>\n\n" + m.content + "<|eot_id|>";
        if (gen) p += "<|start_header_id|>assistant<|end_header_id|>\n\n";
    } else if (tmpl_kind == "zephyr") {
        for (auto& m : msgs) p += "<|" + m.role + "|>\n" + m.content + eos + "\n";
        if (gen) p += "<|assistant|>\n";
    } else if (tmpl_kind == "phi3") {
        for (auto& m : msgs) p += "<|" + m.role + "|>\n" + m.content + "<|end|>\n";
        if (gen) p += "<|assistant|>\n";
    } else if (tmpl_kind == "llama2") {
        std::string sys;
        for (auto& m : msgs) {
            if (m.role == "system") sys = "<<SYS>>\n" + m.content + "\n<</SYS>>\n\n";
            else if (m.role == "user") { p += "[INST] " + sys + m.content + " [/INST]"; sys.clear(); }
            else p += " " + m.content + eos;
        }
    } else {
        for (auto& m : msgs) p += m.role + ": " + m.content + "\n";
        if (gen) p += "assistant:";
    }
    return p;
}

}  // namespace b2srv
