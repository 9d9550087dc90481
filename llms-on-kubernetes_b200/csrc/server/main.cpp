// llama-server -- drop-in for the process the ramalama chart launches:
//     command: ["llama-server"]  args: --host 0.0.0.0 --port 8080 --model <path> --alias <name>
//     (ramalama-models/helm-chart/templates/model-deployments.yaml:26-35)
// and for the HTTP surface its callers use: the Python API gateway's proxy_request
// (ramalama-models/helm-chart/templates/api-gateway.yaml:63-104: urllib, HTTP/1.1 + Connection: close, whole
// response buffered, 300 s timeout), nginx proxy_pass in the vLLM chart's gateway (HTTP/1.0 upstream,
// vllm-models/helm-chart/templates/model-gateway.yaml:51-82), kubelet probes on /health and humans with curl
// (ramalama-models/README.md:273-285).  All compute goes through the C ABI in include/b2.h; without a CUDA
// device the process exits non-zero (no CPU fallback) so that kubelet restarts the pod.
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <signal.h>
#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <functional>
#include <mutex>
#include <random>
#include <fstream>
#include <sstream>
#include <thread>

#include "engine.h"
#include "json.h"
#include "tokenizer.h"

using b2json::Value;
using b2srv::Engine;
using b2srv::GenParams;
using b2srv::GenResult;

namespace {

Engine g_engine;

// ------------------------------------------------------------------------------------------------
// HTTP
// ------------------------------------------------------------------------------------------------
struct Request {
    std::string method, path, version;
    std::vector<std::pair<std::string, std::string>> headers;
    std::string body;
    const std::string* header(const char* name) const {
        for (auto& h : headers)
            if (!strcasecmp(h.first.c_str(), name)) return &h.second;
        return nullptr;
    }
};

bool send_all(int fd, const char* p, size_t n) {
    while (n) {
        ssize_t w = ::send(fd, p, n, MSG_NOSIGNAL);
        if (w <= 0) return false;
        p += w;
        n -= (size_t)w;
    }
    return true;
}
bool send_all(int fd, const std::string& s) { return send_all(fd, s.data(), s.size()); }

const char* reason(int status) {
    switch (status) {
        case 200: return "OK";
        case 204: return "No Content";
        case 400: return "Bad Request";
        case 404: return "Not Found";
        case 405: return "Method Not Allowed";
        case 413: return "Payload Too Large";
        case 500: return "Internal Server Error";
        case 503: return "Service Unavailable";
    }
    return "Error";
}

bool respond(int fd, const Request& rq, int status, const std::string& ctype, const std::string& body, bool keep_alive) {
    std::ostringstream h;
    // answer in the client's protocol version: nginx's default upstream protocol is HTTP/1.0
    h << (rq.version == "HTTP/1.0" ? "HTTP/1.0 " : "HTTP/1.1 ") << status << ' ' << reason(status) << "\r\n"
      << "Content-Type: " << ctype << "\r\n"
      << "Content-Length: " << body.size() << "\r\n"
      << "Access-Control-Allow-Origin: *\r\n"
      << "Server: b2-llama-server\r\n"
      << "Connection: " << (keep_alive ? "keep-alive" : "close") << "\r\n\r\n";
    return send_all(fd, h.str()) && send_all(fd, body);
}

std::string error_json(int status, const std::string& msg, const char* type) {
    Value e = Value::object(), inner = Value::object();
    inner.set("message", msg);
    inner.set("type", type);
    inner.set("code", status);
    e.set("error", inner);
    return b2json::dump(e);
}

// read one request; returns 0 ok, 1 clean close, -1 malformed, -2 too large
int read_request(int fd, std::string& buf, Request& rq) {
    size_t hdr_end;
    while ((hdr_end = buf.find("\r\n\r\n")) == std::string::npos) {
        if (buf.size() > 64 * 1024) return -2;
        char tmp[8192];
        ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
        if (n <= 0) return buf.empty() ? 1 : -1;
        buf.append(tmp, (size_t)n);
    }
    std::istringstream hs(buf.substr(0, hdr_end));
    std::string line;
    if (!std::getline(hs, line)) return -1;
    if (!line.empty() && line.back() == '\r') line.pop_back();
    {
        std::istringstream ls(line);
        if (!(ls >> rq.method >> rq.path >> rq.version)) return -1;
    }
    rq.headers.clear();
    while (std::getline(hs, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        const size_t c = line.find(':');
        if (c == std::string::npos) continue;
        size_t v = c + 1;
        while (v < line.size() && (line[v] == ' ' || line[v] == '\t')) v++;
        rq.headers.emplace_back(line.substr(0, c), line.substr(v));
    }
    size_t clen = 0;
    if (auto* h = rq.header("Content-Length")) clen = (size_t)strtoull(h->c_str(), nullptr, 10);
    if (clen > 64u * 1024 * 1024) return -2;
    if (auto* e = rq.header("Expect"))
        if (!strcasecmp(e->c_str(), "100-continue")) send_all(fd, "HTTP/1.1 100 Continue\r\n\r\n");
    buf.erase(0, hdr_end + 4);
    while (buf.size() < clen) {
        char tmp[65536];
        ssize_t n = ::recv(fd, tmp, sizeof tmp, 0);
        if (n <= 0) return -1;
        buf.append(tmp, (size_t)n);
    }
    rq.body = buf.substr(0, clen);
    buf.erase(0, clen);
    return 0;
}

int64_t unix_now() { return (int64_t)std::chrono::duration_cast<std::chrono::seconds>(std::chrono::system_clock::now().time_since_epoch()).count(); }

std::string new_id(const char* prefix) {
    static std::atomic<uint64_t> ctr{0};
    char b[64];
    snprintf(b, sizeof b, "%s-%llx%04llx", prefix, (unsigned long long)unix_now(), (unsigned long long)(ctr++ & 0xffff));
    return b;
}

void fill_params(const Value& body, GenParams& p) {
    if (body["max_tokens"].is_number()) p.max_tokens = (int)body["max_tokens"].num;
    else if (body["max_completion_tokens"].is_number()) p.max_tokens = (int)body["max_completion_tokens"].num;
    else if (body["n_predict"].is_number()) p.max_tokens = (int)body["n_predict"].num;
    p.temperature = (float)body["temperature"].number_or(p.temperature);
    p.top_p = (float)body["top_p"].number_or(p.top_p);
    p.min_p = (float)body["min_p"].number_or(p.min_p);
    p.top_k = (int)body["top_k"].int_or(p.top_k);
    if (body["seed"].is_number() && body["seed"].num >= 0) p.seed = (uint64_t)body["seed"].num;
    p.ignore_eos = body["ignore_eos"].bool_or(false);
    const Value& st = body["stop"];
    if (st.is_string()) p.stop.push_back(st.str);
    else if (st.is_array())
        for (size_t i = 0; i < st.size(); i++)
            if (st.at(i).is_string()) p.stop.push_back(st.at(i).str);
}

Value usage_json(const GenResult& r) {
    Value u = Value::object();
    u.set("prompt_tokens", r.prompt_tokens);
    u.set("completion_tokens", (int)r.tokens.size());
    u.set("total_tokens", r.prompt_tokens + (int)r.tokens.size());
    return u;
}

// chat and text completions share everything but the prompt construction and the object names
bool handle_completion(int fd, const Request& rq, bool chat, bool keep_alive) {
    Value body;
    std::string perr;
    if (!b2json::parse(rq.body, body, &perr) || !body.is_object())
        return respond(fd, rq, 400, "application/json", error_json(400, "request body is not valid JSON: " + perr, "invalid_request_error"), keep_alive);
    Engine& E = g_engine;
    std::vector<int32_t> prompt;
    if (chat) {
        const Value& msgs = body["messages"];
        if (!msgs.is_array() || msgs.size() == 0)
            return respond(fd, rq, 400, "application/json", error_json(400, "'messages' must be a non-empty array", "invalid_request_error"), keep_alive);
        std::vector<b2srv::ChatMessage> cm;
        for (size_t i = 0; i < msgs.size(); i++) {
            const Value& m = msgs.at(i);
            std::string content;
            if (m["content"].is_string()) content = m["content"].str;
            else if (m["content"].is_array())   // OpenAI content parts: concatenate the text parts
                for (size_t j = 0; j < m["content"].size(); j++) content += m["content"].at(j)["text"].string_or("");
            cm.push_back({m["role"].string_or("user"), content});
        }
        prompt = E.tok.encode(E.tok.apply_chat_template(cm, true), true, true);
    } else {
        const Value& pr = body["prompt"];
        if (pr.is_string()) prompt = E.tok.encode(pr.str, true, false);
        else if (pr.is_array()) {
            // an array of token ids (what the parity harness sends) or of strings (first one is used)
            if (pr.size() && pr.at(0).is_number())
                for (size_t i = 0; i < pr.size(); i++) prompt.push_back((int32_t)pr.at(i).int_or(-1));
            else if (pr.size() && pr.at(0).is_string()) prompt = E.tok.encode(pr.at(0).str, true, false);
        }
        if (prompt.empty())
            return respond(fd, rq, 400, "application/json", error_json(400, "'prompt' must be a string or an array of token ids", "invalid_request_error"), keep_alive);
    }
    GenParams p;
    fill_params(body, p);
    const bool stream = body["stream"].bool_or(false);
    const std::string id = new_id(chat ? "chatcmpl" : "cmpl");
    const int64_t created = unix_now();
    const char* obj = chat ? "chat.completion" : "text_completion";

    if (!stream) {
        GenResult r = E.generate(prompt, p, nullptr);
        if (r.error_status)
            return respond(fd, rq, r.error_status, "application/json",
                           error_json(r.error_status, r.error, r.error_status == 400 ? "invalid_request_error" : "server_error"), keep_alive);
        Value out = Value::object(), choice = Value::object();
        out.set("id", id);
        out.set("object", obj);
        out.set("created", created);
        out.set("model", E.alias);   // the alias the gateway routes on, whatever the request said
        choice.set("index", 0);
        if (chat) {
            Value msg = Value::object();
            msg.set("role", "assistant");
            msg.set("content", r.text);
            choice.set("message", msg);
        } else {
            choice.set("text", r.text);
            Value toks = Value::array();
            for (int32_t t : r.tokens) toks.push(Value((int)t));
            choice.set("tokens", toks);   // extension: generated ids, for id-level parity checks
        }
        choice.set("finish_reason", r.finish);
        Value choices = Value::array();
        choices.push(choice);
        out.set("choices", choices);
        out.set("usage", usage_json(r));
        return respond(fd, rq, 200, "application/json", b2json::dump(out), keep_alive);
    }

    // ---- server-sent events.  Chunked only for 1.1 clients; a 1.0 client gets the stream delimited by close.
    const bool http10 = rq.version == "HTTP/1.0";
    std::string head = std::string(http10 ? "HTTP/1.0" : "HTTP/1.1") + " 200 OK\r\nContent-Type: text/event-stream\r\nCache-Control: no-cache\r\n"
                       "Access-Control-Allow-Origin: *\r\nServer: b2-llama-server\r\n";
    head += http10 ? "Connection: close\r\n\r\n" : "Transfer-Encoding: chunked\r\nConnection: close\r\n\r\n";
    if (!send_all(fd, head)) return false;
    auto send_event = [&](const std::string& data) -> bool {
        const std::string ev = "data: " + data + "\n\n";
        if (http10) return send_all(fd, ev);
        char len[32];
        snprintf(len, sizeof len, "%zx\r\n", ev.size());
        return send_all(fd, len) && send_all(fd, ev) && send_all(fd, "\r\n");
    };
    auto chunk_json = [&](const std::string* piece, const char* finish, bool first) {
        Value out = Value::object(), choice = Value::object();
        out.set("id", id);
        out.set("object", chat ? "chat.completion.chunk" : "text_completion");
        out.set("created", created);
        out.set("model", E.alias);
        choice.set("index", 0);
        if (chat) {
            Value delta = Value::object();
            if (first) delta.set("role", "assistant");
            if (piece) delta.set("content", *piece);
            choice.set("delta", delta);
        } else {
            choice.set("text", piece ? *piece : std::string());
        }
        choice.set("finish_reason", finish ? Value(finish) : Value());
        Value choices = Value::array();
        choices.push(choice);
        out.set("choices", choices);
        return b2json::dump(out);
    };
    bool first = true;
    if (chat) {
        if (!send_event(chunk_json(nullptr, nullptr, true))) return false;
        first = false;
    }
    GenResult r = E.generate(prompt, p, [&](const std::string& piece) { return send_event(chunk_json(&piece, nullptr, false)); });
    if (r.error_status) send_event(error_json(r.error_status, r.error, "server_error"));
    else {
        Value fin;
        b2json::parse(chunk_json(nullptr, r.finish.c_str(), false), fin);
        fin.set("usage", usage_json(r));
        send_event(b2json::dump(fin));
    }
    send_event("[DONE]");
    if (!http10) send_all(fd, "0\r\n\r\n");
    (void)first;
    return false;   // streamed responses close the connection
}

bool handle(int fd, const Request& rq, bool keep_alive) {
    std::string path = rq.path.substr(0, rq.path.find('?'));
    Engine& E = g_engine;
    if (rq.method == "OPTIONS") return respond(fd, rq, 204, "text/plain", "", keep_alive);
    if (path == "/health" || path == "/v1/health") {
        if (rq.method != "GET" && rq.method != "HEAD") return respond(fd, rq, 405, "application/json", error_json(405, "use GET", "invalid_request_error"), keep_alive);
        if (!E.ready) return respond(fd, rq, 503, "application/json", error_json(503, "Loading model", "unavailable_error"), keep_alive);
        return respond(fd, rq, 200, "application/json", "{\"status\":\"ok\"}", keep_alive);
    }
    if (!E.ready) return respond(fd, rq, 503, "application/json", error_json(503, "Loading model", "unavailable_error"), keep_alive);
    if (path == "/v1/models" || path == "/models") {
        if (rq.method != "GET") return respond(fd, rq, 405, "application/json", error_json(405, "use GET", "invalid_request_error"), keep_alive);
        Value list = Value::object(), data = Value::array(), m = Value::object(), meta = Value::object();
        m.set("id", E.alias);
        m.set("object", "model");
        m.set("created", unix_now());
        m.set("owned_by", "llamacpp");
        meta.set("n_vocab", E.info.vocab);
        meta.set("n_ctx_train", E.info.n_ctx_train);
        meta.set("n_embd", E.info.d_model);
        meta.set("size", (uint64_t)E.info.weight_bytes);
        m.set("meta", meta);
        data.push(m);
        list.set("object", "list");
        list.set("data", data);
        return respond(fd, rq, 200, "application/json", b2json::dump(list), keep_alive);
    }
    if (path == "/metrics") {
        // Prometheus text format; llama.cpp's server exports the same names behind --metrics
        if (rq.method != "GET") return respond(fd, rq, 405, "application/json", error_json(405, "use GET", "invalid_request_error"), keep_alive);
        char buf[3072];
        const unsigned long long passes = E.m_passes.load(), cols = E.m_pass_cols.load();
        snprintf(buf, sizeof buf,
                 "# HELP llamacpp:prompt_tokens_total Number of prompt tokens processed.\n# TYPE llamacpp:prompt_tokens_total counter\n"
                 "llamacpp:prompt_tokens_total %llu\n"
                 "# HELP llamacpp:tokens_predicted_total Number of generation tokens processed.\n# TYPE llamacpp:tokens_predicted_total counter\n"
                 "llamacpp:tokens_predicted_total %llu\n"
                 "# HELP llamacpp:requests_processing Number of requests processing.\n# TYPE llamacpp:requests_processing gauge\n"
                 "llamacpp:requests_processing %d\n"
                 "# HELP llamacpp:requests_deferred Number of requests deferred.\n# TYPE llamacpp:requests_deferred gauge\n"
                 "llamacpp:requests_deferred %d\n"
                 "# HELP b2:requests_total Completion requests accepted.\n# TYPE b2:requests_total counter\nb2:requests_total %llu\n"
                 "# HELP b2:prompt_tokens_reused_total Prompt tokens served from a slot's cached prefix.\n# TYPE b2:prompt_tokens_reused_total counter\n"
                 "b2:prompt_tokens_reused_total %llu\n"
                 "# HELP b2:batch_passes_total Batched passes over the weights.\n# TYPE b2:batch_passes_total counter\nb2:batch_passes_total %llu\n"
                 "# HELP b2:batch_columns_total Columns (tokens) in those passes.\n# TYPE b2:batch_columns_total counter\nb2:batch_columns_total %llu\n"
                 "# HELP b2:slots Sequence slots of the context.\n# TYPE b2:slots gauge\nb2:slots %d\n"
                 "# HELP b2:max_batch_columns Widest batch per pass.\n# TYPE b2:max_batch_columns gauge\nb2:max_batch_columns %d\n"
                 "# HELP b2:device_sampled_requests_total Requests whose tokens are drawn on the GPU.\n# TYPE b2:device_sampled_requests_total counter\n"
                 "b2:device_sampled_requests_total %llu\n"
                 "# HELP b2:host_sampled_tokens_total Tokens drawn on the host from a copied logits row.\n# TYPE b2:host_sampled_tokens_total counter\n"
                 "b2:host_sampled_tokens_total %llu\n",
                 (unsigned long long)E.m_prompt_tokens.load(), (unsigned long long)E.m_predicted.load(), E.m_processing.load(), E.m_deferred.load(),
                 (unsigned long long)E.m_requests.load(), (unsigned long long)E.m_prompt_reused.load(), passes, cols, E.n_slots, E.max_cols,
                 (unsigned long long)E.m_dev_sampler_requests.load(), (unsigned long long)E.m_host_sampled.load());
        return respond(fd, rq, 200, "text/plain; version=0.0.4", buf, keep_alive);
    }
    const bool chat = path == "/v1/chat/completions" || path == "/chat/completions";
    const bool text = path == "/v1/completions" || path == "/completions";
    if (chat || text) {
        if (rq.method != "POST") return respond(fd, rq, 405, "application/json", error_json(405, "use POST", "invalid_request_error"), keep_alive);
        return handle_completion(fd, rq, chat, keep_alive);
    }
    if ((path == "/tokenize" || path == "/detokenize") && rq.method == "POST") {
        Value body, out = Value::object();
        if (!b2json::parse(rq.body, body) || !body.is_object())
            return respond(fd, rq, 400, "application/json", error_json(400, "request body is not valid JSON", "invalid_request_error"), keep_alive);
        if (path == "/tokenize") {
            Value toks = Value::array();
            for (int32_t t : E.tok.encode(body["content"].string_or(""), body["add_special"].bool_or(false), body["parse_special"].bool_or(true))) toks.push(Value((int)t));
            out.set("tokens", toks);
        } else {
            std::vector<int32_t> ids;
            for (size_t i = 0; i < body["tokens"].size(); i++) ids.push_back((int32_t)body["tokens"].at(i).int_or(0));
            out.set("content", E.tok.decode(ids));
        }
        return respond(fd, rq, 200, "application/json", b2json::dump(out), keep_alive);
    }
    return respond(fd, rq, 404, "application/json", error_json(404, "File Not Found", "not_found_error"), keep_alive);
}

void serve_connection(int fd) {
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
    std::string buf;
    for (;;) {
        Request rq;
        const int rc = read_request(fd, buf, rq);
        if (rc == 1) break;
        if (rc < 0) {
            Request dummy;
            dummy.version = "HTTP/1.1";
            respond(fd, dummy, rc == -2 ? 413 : 400, "application/json",
                    error_json(rc == -2 ? 413 : 400, rc == -2 ? "request too large" : "malformed HTTP request", "invalid_request_error"), false);
            break;
        }
        bool keep = rq.version == "HTTP/1.1";
        if (auto* c = rq.header("Connection")) {
            if (!strcasecmp(c->c_str(), "close")) keep = false;
            else if (!strcasecmp(c->c_str(), "keep-alive")) keep = true;
        }
        if (!handle(fd, rq, keep) || !keep) break;
    }
    ::shutdown(fd, SHUT_RDWR);
    ::close(fd);
}

void usage() {
    fprintf(stderr,
            "usage: llama-server --model PATH [--alias NAME] [--host IP] [--port N] [--ctx-size N]\n"
            "B200-native GGUF server (OpenAI-compatible /v1/chat/completions, /v1/completions, /v1/models, /health).\n"
            "Other llama-server flags are accepted and ignored.\n");
}

}  // namespace

int main(int argc, char** argv) {
    signal(SIGPIPE, SIG_IGN);
    std::string host = "127.0.0.1", model, alias, tool_mode, tool_arg;
    int port = 8080, n_ctx = 0;
    // sequences decoded together (llama-server's --parallel); the chart cannot pass it, so the default serves concurrency
    int n_parallel = getenv("B2_N_SLOTS") ? atoi(getenv("B2_N_SLOTS")) : 32;
    // flags of upstream llama-server that take a value: accepted and ignored so existing launch lines keep working
    const char* valued[] = {"-t", "--threads", "-tb", "--threads-batch", "-ngl", "--n-gpu-layers", "--gpu-layers", "-b", "--batch-size", "-ub", "--ubatch-size",
                            "-np", "--parallel", "--temp", "--top-k", "--top-p", "--min-p", "-n", "--n-predict", "--chat-template", "-fa", "--flash-attn",
                            "--cache-type-k", "--cache-type-v", "-ctk", "-ctv", "--api-key", "--timeout", "-to", "--threads-http", "--seed", "-s",
                            "--rope-freq-base", "--rope-freq-scale", "--split-mode", "-sm", "--main-gpu", "-mg", "--tensor-split", "-ts", "--log-file",
                            "--slot-save-path", "--lora", "--mmproj", "--device", "-dev", "--numa", "--pooling", "--keep", "--defrag-thold", "-dt"};
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        auto need = [&](const char* what) -> const char* {
            if (i + 1 >= argc) { fprintf(stderr, "llama-server: %s needs a value\n", what); exit(2); }
            return argv[++i];
        };
        if (a == "--host") host = need("--host");
        else if (a == "--port") port = atoi(need("--port"));
        else if (a == "--model" || a == "-m") model = need("--model");
        else if (a == "--alias" || a == "-a") alias = need("--alias");
        else if (a == "--ctx-size" || a == "-c") n_ctx = atoi(need("--ctx-size"));
        else if (a == "--parallel" || a == "-np") n_parallel = atoi(need("--parallel"));
        else if (a == "--tokenize") { tool_mode = "tokenize"; tool_arg = need("--tokenize"); }
        else if (a == "--detokenize") { tool_mode = "detokenize"; tool_arg = need("--detokenize"); }
        else if (a == "--apply-template") { tool_mode = "template"; tool_arg = need("--apply-template"); }
        else if (a == "--pre-split") {
            // offline check of the BPE pre-tokeniser: --pre-split <llama3|qwen2|gpt2> <text> prints the pieces (no model)
            const std::string kind = need("--pre-split");
            const std::string text = need("--pre-split");
            Value out = Value::object(), pieces = Value::array();
            if (!text.empty() && text[0] == '@') {
                // @file: a JSON array of strings -> an array of piece lists
                std::ifstream f(text.substr(1));
                std::stringstream buf;
                buf << f.rdbuf();
                Value arr;
                if (!f || !b2json::parse(buf.str(), arr) || !arr.is_array()) { fprintf(stderr, "llama-server: --pre-split @file needs a JSON array of strings\n"); return 2; }
                for (size_t k = 0; k < arr.size(); k++) {
                    Value one = Value::array();
                    for (const std::string& p : b2srv::bpe_pre_split(kind, arr.at(k).string_or(""))) one.push(Value(p));
                    pieces.push(one);
                }
            } else {
                for (const std::string& p : b2srv::bpe_pre_split(kind, text)) pieces.push(Value(p));
            }
            out.set("pieces", pieces);
            printf("%s\n", b2json::dump(out).c_str());
            return 0;
        }
        else if (a == "--help" || a == "-h") { usage(); return 0; }
        else if (a == "--version") { printf("b2 llama-server (ABI %d)\n", b2_abi_version()); return 0; }
        else {
            bool takes = false;
            for (const char* v : valued) takes |= a == v;
            if (takes && i + 1 < argc) i++;
            fprintf(stderr, "llama-server: ignoring unsupported option %s\n", a.c_str());
        }
    }
    if (model.empty()) { usage(); return 2; }
    if (!tool_mode.empty()) {
        // offline tokenizer tools (no GPU needed): print one JSON document and exit
        b2_gguf* g = nullptr;
        if (b2_gguf_open(model.c_str(), &g) != 0) { fprintf(stderr, "llama-server: %s\n", b2_last_error()); return 1; }
        b2srv::Tokenizer tk;
        std::string err;
        if (!tk.load(g, &err)) { fprintf(stderr, "llama-server: %s\n", err.c_str()); return 1; }
        Value out = Value::object();
        if (tool_mode == "tokenize" || tool_mode == "template") {
            std::string text = tool_arg;
            if (tool_mode == "template") {
                Value msgs;
                if (!b2json::parse(tool_arg, msgs) || !msgs.is_array()) { fprintf(stderr, "llama-server: --apply-template needs a JSON array of messages\n"); return 2; }
                std::vector<b2srv::ChatMessage> cm;
                for (size_t i = 0; i < msgs.size(); i++) cm.push_back({msgs.at(i)["role"].string_or("user"), msgs.at(i)["content"].string_or("")});
                text = tk.apply_chat_template(cm, true);
                out.set("prompt", text);
                out.set("template", tk.template_kind());
            }
            Value ids = Value::array();
            for (int32_t t : tk.encode(text, true, tool_mode == "template")) ids.push(Value((int)t));
            out.set("tokens", ids);
        } else {
            std::vector<int32_t> ids;
            std::stringstream ss(tool_arg);
            std::string item;
            while (std::getline(ss, item, ',')) ids.push_back(atoi(item.c_str()));
            out.set("content", tk.decode(ids));
        }
        printf("%s\n", b2json::dump(out).c_str());
        b2_gguf_close(g);
        return 0;
    }
    if (alias.empty()) alias = model.substr(model.find_last_of('/') == std::string::npos ? 0 : model.find_last_of('/') + 1);
    g_engine.alias = alias;
    g_engine.model_path = model;
    if (b2_device_count() < 1) {
        fprintf(stderr, "llama-server: no CUDA device visible (%s); this server has no CPU fallback\n", b2_last_error());
        return 1;
    }

    const int lfd = ::socket(AF_INET, SOCK_STREAM, 0);
    int one = 1;
    setsockopt(lfd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in addr{};
    addr.sin_family = AF_INET;
    addr.sin_port = htons((uint16_t)port);
    if (inet_pton(AF_INET, host.c_str(), &addr.sin_addr) != 1) { fprintf(stderr, "llama-server: bad --host %s\n", host.c_str()); return 2; }
    if (::bind(lfd, (sockaddr*)&addr, sizeof addr) != 0 || ::listen(lfd, 512) != 0) {
        fprintf(stderr, "llama-server: cannot listen on %s:%d: %s\n", host.c_str(), port, strerror(errno));
        return 1;
    }
    fprintf(stderr, "llama-server: listening on http://%s:%d, loading %s as '%s'\n", host.c_str(), port, model.c_str(), alias.c_str());

    // serve /health (503) while the model loads, like upstream
    std::thread loader([&] {
        const auto t0 = std::chrono::steady_clock::now();
        if (!g_engine.load(model, n_ctx, n_parallel)) {
            fprintf(stderr, "llama-server: failed to load model: %s\n", g_engine.load_error.c_str());
            _exit(1);
        }
        const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        fprintf(stderr, "llama-server: model loaded in %.1f s (%.2f GB in HBM, n_ctx %d, tokenizer %s, template %s); server is ready\n", s,
                g_engine.info.weight_bytes / 1e9, g_engine.n_ctx(), g_engine.tok.model_kind().c_str(), g_engine.tok.template_kind().c_str());
    });
    loader.detach();
    for (;;) {
        const int fd = ::accept(lfd, nullptr, nullptr);
        if (fd < 0) {
            if (errno == EINTR) continue;
            break;
        }
        std::thread(serve_connection, fd).detach();
    }
    return 0;
}
