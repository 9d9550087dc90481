// tokenizer.h -- text <-> token ids from GGUF metadata (the tokenizer + chat template
// stage of the llama-server the chart launches; keys gguf-py constants.py:243-269).
#pragma once
#include <cstdint>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/b2.h"

namespace b2srv {

struct ChatMessage {
    std::string role, content;
};

// the pieces a byte-level BPE model's pre-tokeniser cuts `text` into; kind: "llama3", "qwen2", anything else = gpt-2
std::vector<std::string> bpe_pre_split(const std::string& kind, const std::string& text);

class Tokenizer {
   public:
    // reads tokenizer.ggml.* from the parsed GGUF; returns false (with err) if the vocabulary is missing
    bool load(const b2_gguf* g, std::string* err);

    // parse_special: control / user-defined token strings in the text become their ids
    std::vector<int32_t> encode(const std::string& text, bool add_bos, bool parse_special) const;
    // appends the text of one token to out (byte tokens append one raw byte; control tokens nothing)
    void decode_append(int32_t id, std::string& out) const;
    std::string decode(const std::vector<int32_t>& ids) const;

    // messages[] -> prompt string, using the family the GGUF's tokenizer.chat_template belongs to
    std::string apply_chat_template(const std::vector<ChatMessage>& msgs, bool add_generation_prompt) const;

    bool is_eog(int32_t id) const { return id == eos_id || id == eot_id || (id >= 0 && eog_extra.count(id)); }
    int32_t vocab_size() const { return (int32_t)tokens.size(); }
    const std::string& model_kind() const { return model; }
    const std::string& template_kind() const { return tmpl_kind; }

    int32_t bos_id = -1, eos_id = -1, eot_id = -1, unk_id = -1;
    bool add_bos = true, add_space_prefix = true;

   private:
    void encode_spm(const std::string& text, std::vector<int32_t>& out) const;
    void encode_bpe(const std::string& text, std::vector<int32_t>& out) const;
    void bpe_word(const std::string& word, std::vector<int32_t>& out) const;

    std::string model;       // "llama" (SPM) or "gpt2" (byte-level BPE)
    std::string pre_kind;    // BPE pre-tokeniser family from tokenizer.ggml.pre: "llama3", "qwen2" or "gpt2"
    std::string tmpl, tmpl_kind;
    std::vector<std::string> tokens;
    std::vector<float> scores;
    std::vector<int32_t> types;
    std::unordered_map<std::string, int32_t> tok2id;
    std::unordered_map<std::string, int32_t> merge_rank;   // "left right" -> rank (BPE)
    std::vector<std::pair<std::string, int32_t>> specials;  // longest first
    std::map<int32_t, bool> eog_extra;
    int32_t byte_tok[256];
};

}  // namespace b2srv
