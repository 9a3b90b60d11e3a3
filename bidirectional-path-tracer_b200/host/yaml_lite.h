// Minimal YAML reader for the scene schema the reference's loader accepts.
// The reference parses with yaml-cpp 0.8.0 (src/scene/scene_loader.h:8), which is not in this image;
// this covers the subset its two scenes and its loader's schema use: block mappings, block sequences
// (including "- key: value" items), flow sequences that may nest and span lines, flow mappings,
// single/double quoted scalars, '#' comments (any bytes, incl. non-ASCII) and blank lines.
// Not supported (throws): anchors/aliases, tags, block scalars (| >), multi-document streams.
#pragma once
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace yaml_lite {

struct node {
    enum kind_t { Null, Scalar, Sequence, Map } kind = Null;
    std::string scalar;
    bool quoted = false;
    std::vector<node> seq;
    std::vector<std::pair<std::string, node>> map;   // insertion order kept (material order = file order)

    bool defined() const { return kind != Null; }
    bool is_scalar() const { return kind == Scalar; }
    bool is_sequence() const { return kind == Sequence; }
    bool is_map() const { return kind == Map; }
    size_t size() const { return kind == Sequence ? seq.size() : kind == Map ? map.size() : 0; }
    // operator[] on a non-map or a missing key gives a Null node, like yaml-cpp's const lookup
    const node& operator[](const std::string& key) const {
        static const node null_node;
        if (kind != Map) return null_node;
        for (const auto& kv : map) if (kv.first == key) return kv.second;
        return null_node;
    }
    // yaml-cpp's as<double>/as<int> semantics that matter here: whole-token conversion or failure
    bool as_double(double& out) const {
        if (kind != Scalar || scalar.empty()) return false;
        const std::string& s = scalar;
        if (s == ".inf" || s == "+.inf" || s == ".Inf" || s == ".INF") { out = 1.0 / 0.0; return true; }
        if (s == "-.inf" || s == "-.Inf" || s == "-.INF") { out = -1.0 / 0.0; return true; }
        char* end = nullptr;
        const double v = std::strtod(s.c_str(), &end);
        if (end == s.c_str() || *end != '\0') return false;
        out = v;
        return true;
    }
    bool as_int(int& out) const {
        if (kind != Scalar || scalar.empty()) return false;
        char* end = nullptr;
        const long v = std::strtol(scalar.c_str(), &end, 0);
        if (end == scalar.c_str() || *end != '\0') return false;
        out = int(v);
        return true;
    }
};

namespace detail {

struct line_t { int indent; std::string text; int number; };

// strip a trailing comment that is outside quotes and flow context does not matter for '#'
inline std::string strip_comment(const std::string& s) {
    char q = 0;
    for (size_t i = 0; i < s.size(); ++i) {
        const char c = s[i];
        if (q) { if (c == q) q = 0; continue; }
        if (c == '"' || c == '\'') { q = c; continue; }
        if (c == '#' && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) return s.substr(0, i);
    }
    return s;
}
inline std::string rtrim(std::string s) {
    while (!s.empty() && (s.back() == ' ' || s.back() == '\t' || s.back() == '\r')) s.pop_back();
    return s;
}
inline std::string trim(const std::string& s) {
    size_t a = 0;
    while (a < s.size() && (s[a] == ' ' || s[a] == '\t')) ++a;
    return rtrim(s.substr(a));
}
// bracket depth of a text fragment, ignoring quoted parts
inline int flow_balance(const std::string& s) {
    int d = 0; char q = 0;
    for (char c : s) {
        if (q) { if (c == q) q = 0; continue; }
        if (c == '"' || c == '\'') q = c;
        else if (c == '[' || c == '{') ++d;
        else if (c == ']' || c == '}') --d;
    }
    return d;
}

struct flow_parser {
    const std::string& s; size_t p = 0; int line;
    flow_parser(const std::string& str, int ln) : s(str), line(ln) {}
    [[noreturn]] void fail(const std::string& m) const { throw std::runtime_error("yaml: " + m + " (line " + std::to_string(line) + ")"); }
    void ws() { while (p < s.size() && (s[p] == ' ' || s[p] == '\t' || s[p] == '\n')) ++p; }
    node scalar_until(const char* stops) {
        ws();
        node n; n.kind = node::Scalar;
        if (p < s.size() && (s[p] == '"' || s[p] == '\'')) {
            const char q = s[p++];
            std::string out;
            while (p < s.size() && s[p] != q) {
                if (q == '"' && s[p] == '\\' && p + 1 < s.size()) { ++p; out += (s[p] == 'n' ? '\n' : s[p] == 't' ? '\t' : s[p]); ++p; }
                else out += s[p++];
            }
            if (p >= s.size()) fail("unterminated quoted scalar");
            ++p;
            n.scalar = out; n.quoted = true;
            return n;
        }
        const size_t b = p;
        while (p < s.size() && !std::strchr(stops, s[p])) ++p;
        n.scalar = trim(s.substr(b, p - b));
        if (n.scalar == "~" || n.scalar == "null") { n.kind = node::Null; n.scalar.clear(); }
        return n;
    }
    node value() {
        ws();
        if (p >= s.size()) return node();
        if (s[p] == '[') {
            ++p;
            node n; n.kind = node::Sequence;
            for (;;) {
                ws();
                if (p >= s.size()) fail("unterminated flow sequence");
                if (s[p] == ']') { ++p; break; }
                n.seq.push_back(value_in_flow());
                ws();
                if (p < s.size() && s[p] == ',') ++p;
            }
            return n;
        }
        if (s[p] == '{') {
            ++p;
            node n; n.kind = node::Map;
            for (;;) {
                ws();
                if (p >= s.size()) fail("unterminated flow mapping");
                if (s[p] == '}') { ++p; break; }
                node k = scalar_until(":,}");
                ws();
                node v;
                if (p < s.size() && s[p] == ':') { ++p; v = value_in_flow(); }
                n.map.emplace_back(k.scalar, v);
                ws();
                if (p < s.size() && s[p] == ',') ++p;
            }
            return n;
        }
        return scalar_until("");
    }
    node value_in_flow() {
        ws();
        if (p < s.size() && (s[p] == '[' || s[p] == '{')) return value();
        return scalar_until(",]}");
    }
};

struct parser {
    std::vector<line_t> lines; size_t i = 0;
    [[noreturn]] void fail(const std::string& m, int ln) const { throw std::runtime_error("yaml: " + m + " (line " + std::to_string(ln) + ")"); }

    // parse "text" that follows "key:" or "- " on one logical line (flow text already joined)
    node inline_value(const std::string& text, int ln) {
        const std::string t = trim(text);
        if (t.empty()) return node();
        if (t[0] == '&' || t[0] == '*' || t[0] == '!' || t == "|" || t == ">" || t == "|-" || t == ">-")
            fail("unsupported YAML feature '" + t.substr(0, 1) + "'", ln);
        flow_parser fp(t, ln);
        node n = fp.value();
        fp.ws();
        if (fp.p != t.size()) fail("trailing characters after value", ln);
        return n;
    }
    // split "key: rest" at the first ": " / trailing ':' outside quotes and brackets; false if no key
    static bool split_key(const std::string& t, std::string& key, std::string& rest) {
        char q = 0; int d = 0;
        for (size_t k = 0; k < t.size(); ++k) {
            const char c = t[k];
            if (q) { if (c == q) q = 0; continue; }
            if (c == '"' || c == '\'') { q = c; continue; }
            if (c == '[' || c == '{') ++d;
            else if (c == ']' || c == '}') --d;
            else if (c == ':' && d == 0 && (k + 1 == t.size() || t[k + 1] == ' ' || t[k + 1] == '\t')) {
                key = trim(t.substr(0, k));
                if (key.size() >= 2 && (key.front() == '"' || key.front() == '\'') && key.back() == key.front())
                    key = key.substr(1, key.size() - 2);
                rest = t.substr(k + 1);
                return true;
            }
        }
        return false;
    }

    node block(int indent) {
        if (i >= lines.size() || lines[i].indent < indent) return node();
        const int my = lines[i].indent;
        if (lines[i].text[0] == '-' && (lines[i].text.size() == 1 || lines[i].text[1] == ' ')) return sequence(my);
        std::string k, r;
        if (split_key(lines[i].text, k, r)) return mapping(my);
        node n = inline_value(lines[i].text, lines[i].number);
        ++i;
        return n;
    }
    node mapping(int indent) {
        node n; n.kind = node::Map;
        while (i < lines.size() && lines[i].indent == indent) {
            std::string key, rest;
            if (!split_key(lines[i].text, key, rest)) fail("expected 'key: value'", lines[i].number);
            const int ln = lines[i].number;
            ++i;
            node v;
            if (!trim(rest).empty()) v = inline_value(rest, ln);
            else if (i < lines.size() && (lines[i].indent > indent ||
                     (lines[i].indent == indent && lines[i].text[0] == '-' && (lines[i].text.size() == 1 || lines[i].text[1] == ' '))))
                v = block(lines[i].indent);
            n.map.emplace_back(key, v);
        }
        if (i < lines.size() && lines[i].indent > indent) fail("bad indentation", lines[i].number);
        return n;
    }
    node sequence(int indent) {
        node n; n.kind = node::Sequence;
        while (i < lines.size() && lines[i].indent == indent && lines[i].text[0] == '-' &&
               (lines[i].text.size() == 1 || lines[i].text[1] == ' ')) {
            const int ln = lines[i].number;
            std::string rest = lines[i].text.substr(1);
            size_t lead = 0;
            while (lead < rest.size() && rest[lead] == ' ') ++lead;
            rest = rest.substr(lead);
            if (rest.empty()) {
                ++i;
                n.seq.push_back(i < lines.size() && lines[i].indent > indent ? block(lines[i].indent) : node());
                continue;
            }
            std::string k, r;
            if (rest[0] != '[' && rest[0] != '{' && rest[0] != '"' && rest[0] != '\'' && split_key(rest, k, r)) {
                // "- key: value": the item is a mapping whose first key sits at column indent+1+lead
                lines[i].indent = indent + 1 + int(lead);
                lines[i].text = rest;
                n.seq.push_back(mapping(lines[i].indent));
            } else {
                n.seq.push_back(inline_value(rest, ln));
                ++i;
            }
        }
        return n;
    }
};

}  // namespace detail

inline node parse(std::istream& in) {
    detail::parser P;
    std::string raw;
    int number = 0;
    std::string pending; int pending_indent = 0, pending_line = 0, depth = 0;
    while (std::getline(in, raw)) {
        ++number;
        if (number == 1 && raw.size() >= 3 && (unsigned char)raw[0] == 0xEF && (unsigned char)raw[1] == 0xBB && (unsigned char)raw[2] == 0xBF)
            raw = raw.substr(3);
        std::string s = detail::rtrim(detail::strip_comment(raw));
        if (depth > 0) {   // continuation of a multi-line flow collection
            pending += " " + detail::trim(s);
            depth += detail::flow_balance(s);
            if (depth <= 0) { P.lines.push_back({pending_indent, pending, pending_line}); pending.clear(); depth = 0; }
            continue;
        }
        size_t ind = 0;
        while (ind < s.size() && s[ind] == ' ') ++ind;
        if (ind < s.size() && s[ind] == '\t') throw std::runtime_error("yaml: tab indentation (line " + std::to_string(number) + ")");
        if (ind >= s.size()) continue;   // blank or comment-only
        const std::string text = s.substr(ind);
        if (text == "---" && P.lines.empty()) continue;
        if (text == "---" || text == "...") throw std::runtime_error("yaml: multi-document streams are not supported");
        const int bal = detail::flow_balance(text);
        if (bal > 0) { pending = text; pending_indent = int(ind); pending_line = number; depth = bal; continue; }
        P.lines.push_back({int(ind), text, number});
    }
    if (depth > 0) throw std::runtime_error("yaml: unterminated flow collection starting at line " + std::to_string(pending_line));
    if (P.lines.empty()) return node();
    node root = P.block(P.lines[0].indent);
    if (P.i != P.lines.size()) P.fail("unexpected content", P.lines[P.i].number);
    return root;
}

inline node load_file(const std::string& path) {
    std::ifstream in(path);
    if (!in) throw std::runtime_error("bad file: " + path);   // yaml-cpp: YAML::BadFile
    return parse(in);
}

}  // namespace yaml_lite
