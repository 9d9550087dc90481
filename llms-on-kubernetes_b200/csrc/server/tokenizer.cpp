// tokenizer.cpp -- SentencePiece-style ("llama") and byte-level BPE ("gpt2") tokenisation driven by the
// GGUF's tokenizer.ggml.* metadata, plus the chat-template families the reference's example models use
// (Zephyr for TinyLlama, ChatML for Qwen, Llama-3, Llama-2/Mistral [INST]).  Replaces the tokenizer stage
// of the llama-server the chart launches.
//
// Fidelity: the SPM path follows llama.cpp's algorithm (UTF-8 symbols, highest-score bigram merged first,
// <0xXX> byte fallback).  The BPE path applies the GGUF's merge ranks over the pre-split the model's
// tokenizer.ggml.pre names: the llama-3, qwen2 and gpt-2 regexes are matched by hand over Unicode code points
// (bpe_pre_split; letters and numbers from unicode_tables.h), checked piece for piece against the `tokenizers`
// library running the same regexes (tests/test_tokenizer_presplit.py).  Templates are recognised by their markers,
// not interpreted as Jinja.
#include "tokenizer.h"

#include <algorithm>
#include <cstring>
#include <queue>

#include "unicode_tables.h"

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
    {
        // llama.cpp's names for the pre-tokeniser regex families (tokenizer.ggml.pre)
        std::string pre;
        kv_string(g, "tokenizer.ggml.pre", &pre);
        if (pre == "llama-bpe" || pre == "llama3" || pre == "llama-v3") pre_kind = "llama3";
        else if (pre == "qwen2" || pre == "deepseek-r1-qwen") pre_kind = "qwen2";
        else pre_kind = "gpt2";
    }
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
template <size_t NR>
bool in_ranges(const uint32_t (&r)[NR][2], uint32_t cp) {
    size_t lo = 0, hi = NR;
    while (lo < hi) {
        const size_t mid = (lo + hi) / 2;
        if (cp < r[mid][0]) hi = mid;
        else if (cp > r[mid][1]) lo = mid + 1;
        else return true;
    }
    return false;
}
bool cp_letter(uint32_t c) { return c < 0x80 ? ((c | 0x20) >= 'a' && (c | 0x20) <= 'z') : in_ranges(UNICODE_LETTER_RANGES, c); }
bool cp_number(uint32_t c) { return c < 0x80 ? (c >= '0' && c <= '9') : in_ranges(UNICODE_NUMBER_RANGES, c); }
// \s of the regex engines: Unicode White_Space
bool cp_space(uint32_t c) {
    return (c >= 9 && c <= 13) || c == 0x20 || c == 0x85 || c == 0xA0 || c == 0x1680 || (c >= 0x2000 && c <= 0x200A) || c == 0x2028 || c == 0x2029 ||
           c == 0x202F || c == 0x205F || c == 0x3000;
}
bool cp_newline(uint32_t c) { return c == '\r' || c == '\n'; }
bool cp_symbol(uint32_t c) { return !cp_space(c) && !cp_letter(c) && !cp_number(c); }   // [^\s\p{L}\p{N}]
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

// The pre-tokeniser regexes, matched by hand (first alternative that matches wins, as in the regex engines):
//   llama3: (?i:'s|'t|'re|'ve|'m|'ll|'d) | [^\r\n\p{L}\p{N}]?\p{L}+ | \p{N}{1,3} | ?[^\s\p{L}\p{N}]+[\r\n]* | \s*[\r\n]+ | \s+(?!\S) | \s+
//   qwen2 : the same with \p{N} one digit at a time
//   gpt2  : 's|'t|'re|'ve|'m|'ll|'d | ?\p{L}+ | ?\p{N}+ | ?[^\s\p{L}\p{N}]+ | \s+(?!\S) | \s+
std::vector<std::string> bpe_pre_split(const std::string& kind, const std::string& text) {
    const bool gpt2 = kind != "llama3" && kind != "qwen2";
    const int max_digits = kind == "llama3" ? 3 : kind == "qwen2" ? 1 : 1 << 30;
    // code points with their byte offsets (invalid bytes pass through as themselves)
    std::vector<uint32_t> cp;
    std::vector<size_t> off;
    for (size_t i = 0; i < text.size();) {
        const unsigned char c = text[i];
        size_t len = utf8_len(c);
        if (len > text.size() - i) len = 1;
        uint32_t v = c;
        if (len == 2) v = ((c & 0x1F) << 6) | (text[i + 1] & 0x3F);
        else if (len == 3) v = ((c & 0x0F) << 12) | ((text[i + 1] & 0x3F) << 6) | (text[i + 2] & 0x3F);
        else if (len == 4) v = ((c & 0x07) << 18) | ((text[i + 1] & 0x3F) << 12) | ((text[i + 2] & 0x3F) << 6) | (text[i + 3] & 0x3F);
        cp.push_back(v);
        off.push_back(i);
        i += len;
    }
    const size_t n = cp.size();
    off.push_back(text.size());
    auto at = [&](size_t i) -> uint32_t { return i < n ? cp[i] : 0xFFFFFFFFu; };
    auto lower = [](uint32_t c) { return c < 0x80 ? (uint32_t)(c | 0x20) : c; };
    std::vector<std::string> out;
    size_t i = 0;
    while (i < n) {
        size_t j = i;
        const uint32_t c = cp[i];
        // 1. contractions
        if (c == '\'' && i + 1 < n) {
            const uint32_t a = gpt2 ? cp[i + 1] : lower(cp[i + 1]), b = i + 2 < n ? (gpt2 ? cp[i + 2] : lower(cp[i + 2])) : 0;
            if (a == 's' || a == 't' || a == 'm' || a == 'd') j = i + 2;
            else if ((a == 'r' && b == 'e') || (a == 'v' && b == 'e') || (a == 'l' && b == 'l')) j = i + 3;
        }
        if (j == i) {
            if (gpt2) {
                const size_t k = (c == ' ' && i + 1 < n) ? i + 1 : i;   // optional leading space
                if (cp_letter(at(k))) { j = k; while (j < n && cp_letter(cp[j])) j++; }
                else if (cp_number(at(k))) { j = k; while (j < n && cp_number(cp[j])) j++; }
                else if (k < n && cp_symbol(cp[k])) { j = k; while (j < n && cp_symbol(cp[j])) j++; }
            } else {
                // 2. [^\r\n\p{L}\p{N}]?\p{L}+
                if (cp_letter(c)) { j = i; while (j < n && cp_letter(cp[j])) j++; }
                else if (!cp_newline(c) && !cp_number(c) && cp_letter(at(i + 1))) { j = i + 1; while (j < n && cp_letter(cp[j])) j++; }
                // 3. \p{N}{1,3} (qwen2: one)
                else if (cp_number(c)) { int cnt = 0; while (j < n && cp_number(cp[j]) && cnt < max_digits) { j++; cnt++; } }
                else {
                    // 4.  ?[^\s\p{L}\p{N}]+[\r\n]*
                    const size_t k = (c == ' ' && i + 1 < n && cp_symbol(cp[i + 1])) ? i + 1 : i;
                    if (cp_symbol(at(k))) {
                        j = k;
                        while (j < n && cp_symbol(cp[j])) j++;
                        while (j < n && cp_newline(cp[j])) j++;
                    }
                }
            }
        }
        if (j == i) {
            // whitespace run [i, e)
            size_t e = i;
            while (e < n && cp_space(cp[e])) e++;
            if (e == i) e = i + 1;   // (cannot happen: every code point is letter, number, symbol or space)
            j = e;
            size_t last_nl = n;
            if (!gpt2)
                for (size_t k = i; k < e; k++)
                    if (cp_newline(cp[k])) last_nl = k;
            if (last_nl != n) j = last_nl + 1;            // 5. \s*[\r\n]+
            else if (e < n && e - i > 1) j = e - 1;       // 6. \s+(?!\S): leave the last space to the next piece
        }
        out.push_back(text.substr(off[i], off[j] - off[i]));
        i = j;
    }
    return out;
}

void Tokenizer::encode_bpe(const std::string& text, std::vector<int32_t>& out) const {
    for (const std::string& piece : bpe_pre_split(pre_kind, text)) bpe_word(piece, out);
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
