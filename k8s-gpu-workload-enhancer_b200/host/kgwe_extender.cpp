// kgwe_extender.cpp — see kgwe_extender.hpp.
#include "kgwe_extender.hpp"

#include <arpa/inet.h>
#include <netinet/in.h>
#include <sys/socket.h>
#include <unistd.h>

#include <cctype>
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace kgwe_host {

// ---------------------------------------------------------------- JSON
namespace {
struct Parser {
    const std::string& s; size_t i = 0; std::string err;
    explicit Parser(const std::string& t) : s(t) {}
    void ws() { while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\n' || s[i] == '\r')) ++i; }
    bool fail(const char* m) { if (err.empty()) err = std::string(m) + " at offset " + std::to_string(i); return false; }
    static void utf8(std::string& o, unsigned cp) {
        if (cp < 0x80) o += (char)cp;
        else if (cp < 0x800) { o += (char)(0xC0 | (cp >> 6)); o += (char)(0x80 | (cp & 0x3F)); }
        else if (cp < 0x10000) { o += (char)(0xE0 | (cp >> 12)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
        else { o += (char)(0xF0 | (cp >> 18)); o += (char)(0x80 | ((cp >> 12) & 0x3F)); o += (char)(0x80 | ((cp >> 6) & 0x3F)); o += (char)(0x80 | (cp & 0x3F)); }
    }
    bool hex4(unsigned* v) {
        if (i + 4 > s.size()) return fail("short \\u escape");
        unsigned x = 0;
        for (int k = 0; k < 4; ++k) { const char c = s[i++]; x <<= 4; if (c >= '0' && c <= '9') x |= c - '0'; else if (c >= 'a' && c <= 'f') x |= c - 'a' + 10; else if (c >= 'A' && c <= 'F') x |= c - 'A' + 10; else return fail("bad \\u escape"); }
        *v = x; return true;
    }
    bool string(std::string* out) {
        if (i >= s.size() || s[i] != '"') return fail("expected a string");
        ++i; out->clear();
        while (i < s.size() && s[i] != '"') {
            char c = s[i++];
            if (c != '\\') { *out += c; continue; }
            if (i >= s.size()) return fail("short escape");
            c = s[i++];
            switch (c) {
                case '"': *out += '"'; break; case '\\': *out += '\\'; break; case '/': *out += '/'; break;
                case 'b': *out += '\b'; break; case 'f': *out += '\f'; break; case 'n': *out += '\n'; break;
                case 'r': *out += '\r'; break; case 't': *out += '\t'; break;
                case 'u': {
                    unsigned cp = 0; if (!hex4(&cp)) return false;
                    if (cp >= 0xD800 && cp < 0xDC00 && i + 1 < s.size() && s[i] == '\\' && s[i + 1] == 'u') { i += 2; unsigned lo = 0; if (!hex4(&lo)) return false; cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00); }
                    utf8(*out, cp); break;
                }
                default: return fail("bad escape");
            }
        }
        if (i >= s.size()) return fail("unterminated string");
        ++i; return true;
    }
    bool value(Json* v, int depth) {
        if (depth > 64) return fail("nesting too deep");
        ws();
        if (i >= s.size()) return fail("unexpected end");
        const char c = s[i];
        if (c == '{') {
            v->type = Json::Object; ++i; ws();
            if (i < s.size() && s[i] == '}') { ++i; return true; }
            for (;;) {
                ws(); std::string k; if (!string(&k)) return false;
                ws(); if (i >= s.size() || s[i] != ':') return fail("expected ':'");
                ++i; Json child; if (!value(&child, depth + 1)) return false;
                v->fields.emplace_back(std::move(k), std::move(child));
                ws(); if (i < s.size() && s[i] == ',') { ++i; continue; }
                if (i < s.size() && s[i] == '}') { ++i; return true; }
                return fail("expected ',' or '}'");
            }
        }
        if (c == '[') {
            v->type = Json::Array; ++i; ws();
            if (i < s.size() && s[i] == ']') { ++i; return true; }
            for (;;) {
                Json child; if (!value(&child, depth + 1)) return false;
                v->items.push_back(std::move(child));
                ws(); if (i < s.size() && s[i] == ',') { ++i; continue; }
                if (i < s.size() && s[i] == ']') { ++i; return true; }
                return fail("expected ',' or ']'");
            }
        }
        if (c == '"') { v->type = Json::String; return string(&v->str); }
        if (s.compare(i, 4, "true") == 0) { v->type = Json::Bool; v->b = true; i += 4; return true; }
        if (s.compare(i, 5, "false") == 0) { v->type = Json::Bool; v->b = false; i += 5; return true; }
        if (s.compare(i, 4, "null") == 0) { v->type = Json::Null; i += 4; return true; }
        if (c == '-' || (c >= '0' && c <= '9')) {
            const size_t b = i;
            while (i < s.size() && (std::isdigit((unsigned char)s[i]) || s[i] == '-' || s[i] == '+' || s[i] == '.' || s[i] == 'e' || s[i] == 'E')) ++i;
            v->type = Json::Number; v->str = s.substr(b, i - b);
            char* end = nullptr; v->num = strtod(v->str.c_str(), &end);
            if (!end || *end) return fail("bad number");
            return true;
        }
        return fail("unexpected character");
    }
};
bool ieq(const std::string& a, const std::string& b) {
    if (a.size() != b.size()) return false;
    for (size_t k = 0; k < a.size(); ++k) if (std::tolower((unsigned char)a[k]) != std::tolower((unsigned char)b[k])) return false;
    return true;
}
}  // namespace

bool Json::Parse(const std::string& text, Json* out, std::string* err) {
    Parser p(text);
    *out = Json();
    if (!p.value(out, 0)) { if (err) *err = p.err; return false; }
    p.ws();
    if (p.i != text.size()) { if (err) *err = "trailing characters at offset " + std::to_string(p.i); return false; }
    return true;
}

const Json* Json::get(const std::string& key) const {
    if (type != Object) return nullptr;
    for (auto& f : fields) if (f.first == key) return &f.second;          // exact match first, as encoding/json prefers
    for (auto& f : fields) if (ieq(f.first, key)) return &f.second;
    return nullptr;
}

std::string Json::Quote(const std::string& s) {
    std::string o = "\"";
    for (unsigned char c : s) {
        switch (c) {
            case '"': o += "\\\""; break; case '\\': o += "\\\\"; break; case '\n': o += "\\n"; break;
            case '\r': o += "\\r"; break; case '\t': o += "\\t"; break;
            default: if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; } else o += (char)c;
        }
    }
    return o + "\"";
}

// ---------------------------------------------------------------- pod -> workload
namespace {
// A resource quantity as kube-scheduler prints it for extended resources: an integer, as a JSON string or number.
bool quantity(const Json& q, long* out) {
    if (q.type == Json::Number) { *out = (long)q.num; return q.num == std::floor(q.num) && q.num >= 0; }
    if (q.type != Json::String || q.str.empty()) return false;
    char* end = nullptr; const long v = strtol(q.str.c_str(), &end, 10);
    if (!end || *end || v < 0) return false;
    *out = v; return true;
}
std::string str_of(const Json* j) { return j && j->type == Json::String ? j->str : std::string(); }
}  // namespace

std::string WorkloadFromPod(const Json& pod, GPUWorkload* w) {
    if (pod.type != Json::Object) return "pod is not an object";
    *w = GPUWorkload();
    const Json* meta = pod.get("metadata");
    w->Name = str_of(meta ? meta->get("name") : nullptr);
    w->Namespace = str_of(meta ? meta->get("namespace") : nullptr);
    w->UID = str_of(meta ? meta->get("uid") : nullptr);
    if (w->UID.empty()) return "pod has no metadata.uid";
    const Json* spec = pod.get("spec");
    const Json* containers = spec ? spec->get("containers") : nullptr;
    long gpus = 0;
    if (containers && containers->type == Json::Array)
        for (auto& c : containers->items) {
            const Json* res = c.get("resources");
            const Json* lim = res ? res->get("limits") : nullptr;
            if (!lim || lim->type != Json::Object) continue;
            for (auto& f : lim->fields) {
                long q = 0;
                if (f.first == "nvidia.com/gpu") { if (!quantity(f.second, &q)) return "bad quantity for nvidia.com/gpu"; gpus += q; }
                else if (f.first.compare(0, 15, "nvidia.com/mig-") == 0 && !w->Requirements.MIGReq) {
                    if (!quantity(f.second, &q)) return "bad quantity for " + f.first;
                    w->Requirements.MIGReq = std::make_shared<MIGRequirements>();
                    w->Requirements.MIGReq->Profile = f.first.substr(15); w->Requirements.MIGReq->Count = (int)q;
                }
            }
        }
    if (gpus > 1000000) return "nvidia.com/gpu quantity out of range";
    w->Requirements.GPUCount = (int)gpus;
    const Json* ann = meta ? meta->get("annotations") : nullptr;
    if (ann && ann->type == Json::Object) {
        const std::string pref = str_of(ann->get("kgwe.io/topology-preference"));
        if (!pref.empty()) w->Requirements.TopologyPreference = pref;
        const std::string mem = str_of(ann->get("kgwe.io/min-memory-gb"));
        if (!mem.empty()) { char* e = nullptr; const long v = strtol(mem.c_str(), &e, 10); if (!e || *e || v < 0 || v > (1 << 20)) return "bad kgwe.io/min-memory-gb"; w->Requirements.MinMemoryGB = (int)v; }
        w->Requirements.GPUArchitecture = str_of(ann->get("kgwe.io/gpu-architecture"));
        const std::string bw = str_of(ann->get("kgwe.io/min-bandwidth-gbps"));
        if (!bw.empty()) { char* e = nullptr; const double v = strtod(bw.c_str(), &e); if (!e || *e || !(v >= 0)) return "bad kgwe.io/min-bandwidth-gbps"; w->Requirements.MinBandwidthGBps = v; }
    }
    return "";
}

// ---------------------------------------------------------------- verbs
bool ExtenderServer::NodeNamesOf(const Json& args, std::vector<std::string>* names, bool* as_nodes) const {
    names->clear(); *as_nodes = false;
    if (const Json* nn = args.get("nodenames")) {
        if (nn->type == Json::Array) { for (auto& n : nn->items) if (n.type == Json::String) names->push_back(n.str); return true; }
    }
    if (const Json* nodes = args.get("nodes")) {          // nodeCacheCapable: false -> the scheduler sends a v1.NodeList
        const Json* items = nodes->get("items");
        if (items && items->type == Json::Array) {
            *as_nodes = true;
            for (auto& n : items->items) { const Json* m = n.get("metadata"); const std::string name = str_of(m ? m->get("name") : nullptr); if (!name.empty()) names->push_back(name); }
            return true;
        }
    }
    return false;
}

std::string ExtenderServer::HandleFilter(const std::string& body) {
    Json args; std::string err;
    if (!Json::Parse(body, &args, &err)) return "{\"error\":" + Json::Quote("bad ExtenderArgs: " + err) + "}";
    const Json* pod = args.get("pod");
    GPUWorkload w;
    if (!pod || !(err = WorkloadFromPod(*pod, &w)).empty()) return "{\"error\":" + Json::Quote("bad pod: " + (pod ? err : std::string("missing"))) + "}";
    std::vector<std::string> names; bool as_nodes = false;
    if (!NodeNamesOf(args, &names, &as_nodes)) return "{\"error\":\"ExtenderArgs carries neither nodes nor nodenames\"}";
    seen_[w.UID] = w;
    const ExtenderFilterResult r = sched_->Filter(w, names);
    std::string o = "{";
    if (as_nodes) {
        o += "\"nodes\":{\"kind\":\"NodeList\",\"apiVersion\":\"v1\",\"items\":[";
        for (size_t i = 0; i < r.NodeNames.size(); ++i) o += std::string(i ? "," : "") + "{\"metadata\":{\"name\":" + Json::Quote(r.NodeNames[i]) + "}}";
        o += "]},";
    } else {
        o += "\"nodenames\":[";
        for (size_t i = 0; i < r.NodeNames.size(); ++i) o += std::string(i ? "," : "") + Json::Quote(r.NodeNames[i]);
        o += "],";
    }
    o += "\"failedNodes\":{";
    bool first = true;
    for (auto& kv : r.FailedNodes) { o += std::string(first ? "" : ",") + Json::Quote(kv.first) + ":" + Json::Quote(kv.second); first = false; }
    o += "},\"error\":" + Json::Quote(r.Error) + "}";
    return o;
}

std::string ExtenderServer::HandlePrioritize(const std::string& body) {
    Json args; std::string err;
    if (!Json::Parse(body, &args, &err)) return "[]";
    const Json* pod = args.get("pod");
    GPUWorkload w;
    if (!pod || !WorkloadFromPod(*pod, &w).empty()) return "[]";
    std::vector<std::string> names; bool as_nodes = false;
    if (!NodeNamesOf(args, &names, &as_nodes)) return "[]";
    seen_[w.UID] = w;
    const std::vector<HostPriority> pr = sched_->Prioritize(w, names);
    std::string o = "[";
    for (size_t i = 0; i < pr.size(); ++i) o += std::string(i ? "," : "") + "{\"host\":" + Json::Quote(pr[i].Host) + ",\"score\":" + std::to_string(pr[i].Score) + "}";
    return o + "]";
}

std::string ExtenderServer::HandleBind(const std::string& body) {
    Json args; std::string err;
    if (!Json::Parse(body, &args, &err)) return "{\"error\":" + Json::Quote("bad ExtenderBindingArgs: " + err) + "}";
    const std::string uid = str_of(args.get("podUID")), node = str_of(args.get("node"));
    if (uid.empty() || node.empty()) return "{\"error\":\"ExtenderBindingArgs needs podUID and node\"}";
    auto it = seen_.find(uid);
    if (it == seen_.end()) return "{\"error\":" + Json::Quote("pod " + uid + " was never filtered or prioritised here") + "}";
    const Result<SchedulingDecision> d = sched_->Bind(it->second, node);
    if (!d.ok()) return "{\"error\":" + Json::Quote(d.error) + "}";
    seen_.erase(it);
    std::string o = "{\"error\":\"\",\"selectedGPUs\":[";          // (selectedGPUs: an addition the scheduler ignores; the device plugin side can use it)
    for (size_t i = 0; i < d.value.SelectedGPUs.size(); ++i) o += std::string(i ? "," : "") + Json::Quote(d.value.SelectedGPUs[i]);
    return o + "],\"score\":" + std::to_string(d.value.Score) + "}";
}

std::string ExtenderServer::MetricsText() { return sched_->MetricsText(); }

// ---------------------------------------------------------------- HTTP
std::string ExtenderServer::Route(const std::string& method, const std::string& path, const std::string& body, int* status, std::string* ctype) {
    *status = 200; *ctype = "application/json";
    if (method == "POST" && path == "/filter") return HandleFilter(body);
    if (method == "POST" && path == "/prioritize") return HandlePrioritize(body);
    if (method == "POST" && path == "/bind") return HandleBind(body);
    if (method == "GET" && path == "/metrics") { *ctype = "text/plain; version=0.0.4"; return MetricsText(); }
    if (method == "GET" && path == "/healthz") { *ctype = "text/plain"; return "ok\n"; }
    if (method == "POST" && path == "/quitquitquit") { stop_ = true; *ctype = "text/plain"; return "bye\n"; }
    *status = 404; *ctype = "text/plain";
    return "not found\n";
}

std::string ExtenderServer::Serve(int port, void (*on_listen)(int)) {
    listen_fd_ = socket(AF_INET, SOCK_STREAM, 0);
    if (listen_fd_ < 0) return std::string("socket: ") + strerror(errno);
    int one = 1; setsockopt(listen_fd_, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
    sockaddr_in addr; memset(&addr, 0, sizeof addr);
    addr.sin_family = AF_INET; addr.sin_addr.s_addr = htonl(INADDR_ANY); addr.sin_port = htons((uint16_t)port);
    if (bind(listen_fd_, (sockaddr*)&addr, sizeof addr) < 0 || listen(listen_fd_, 64) < 0) { const std::string e = std::string("bind/listen: ") + strerror(errno); close(listen_fd_); return e; }
    socklen_t len = sizeof addr;
    if (on_listen && getsockname(listen_fd_, (sockaddr*)&addr, &len) == 0) on_listen(ntohs(addr.sin_port));
    while (!stop_) {
        const int fd = accept(listen_fd_, nullptr, nullptr);
        if (fd < 0) { if (errno == EINTR) continue; break; }
        std::string req; char buf[8192]; size_t header_end = std::string::npos; size_t need = 0;
        for (;;) {
            const ssize_t n = recv(fd, buf, sizeof buf, 0);
            if (n <= 0) break;
            req.append(buf, (size_t)n);
            if (header_end == std::string::npos) {
                header_end = req.find("\r\n\r\n");
                if (header_end != std::string::npos) {
                    std::string head = req.substr(0, header_end);
                    for (auto& c : head) c = (char)std::tolower((unsigned char)c);
                    const size_t cl = head.find("content-length:");
                    need = cl == std::string::npos ? 0 : (size_t)strtoul(head.c_str() + cl + 15, nullptr, 10);
                    if (need > (64u << 20)) { header_end = std::string::npos; break; }
                }
            }
            if (header_end != std::string::npos && req.size() >= header_end + 4 + need) break;
            if (req.size() > (65u << 20)) break;
        }
        int status = 400; std::string ctype = "text/plain", out = "bad request\n";
        if (header_end != std::string::npos && req.size() >= header_end + 4 + need) {
            const size_t sp1 = req.find(' '), sp2 = sp1 == std::string::npos ? sp1 : req.find(' ', sp1 + 1);
            if (sp2 != std::string::npos) {
                std::string path = req.substr(sp1 + 1, sp2 - sp1 - 1);
                const size_t q = path.find('?'); if (q != std::string::npos) path.resize(q);
                out = Route(req.substr(0, sp1), path, req.substr(header_end + 4, need), &status, &ctype);
            }
        }
        char head[256];
        const int hl = snprintf(head, sizeof head, "HTTP/1.1 %d %s\r\nContent-Type: %s\r\nContent-Length: %zu\r\nConnection: close\r\n\r\n",
                                status, status == 200 ? "OK" : status == 404 ? "Not Found" : "Bad Request", ctype.c_str(), out.size());
        std::string resp(head, (size_t)hl); resp += out;
        size_t off = 0;
        while (off < resp.size()) { const ssize_t n = send(fd, resp.data() + off, resp.size() - off, MSG_NOSIGNAL); if (n <= 0) break; off += (size_t)n; }
        close(fd);
    }
    close(listen_fd_); listen_fd_ = -1;
    return "";
}

}  // namespace kgwe_host
