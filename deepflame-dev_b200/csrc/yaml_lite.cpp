#include "yaml_lite.hpp"

#include <cstdlib>
#include <fstream>
#include <sstream>

namespace dfchem {
namespace yaml {

bool Node::has(const std::string& key) const {
    for (auto& kv : map)
        if (kv.first == key) return true;
    return false;
}
const Node& Node::at(const std::string& key) const {
    for (auto& kv : map)
        if (kv.first == key) return *kv.second;
    throw std::runtime_error("yaml: missing key '" + key + "'");
}
const Node& Node::at(size_t i) const {
    if (kind != Seq || i >= seq.size()) throw std::runtime_error("yaml: bad sequence index");
    return *seq[i];
}
double Node::as_double() const {
    if (kind != Scalar) throw std::runtime_error("yaml: not a scalar");
    char* end = nullptr;
    double v = std::strtod(scalar.c_str(), &end);
    if (end == scalar.c_str()) throw std::runtime_error("yaml: not a number: '" + scalar + "'");
    return v;
}
const std::string& Node::as_string() const {
    if (kind != Scalar) throw std::runtime_error("yaml: not a scalar");
    return scalar;
}

namespace {

struct Line {
    int indent;
    std::string text;  // without indent, without trailing comment/space
    int lineno;
};

std::string rtrim(std::string s) {
    while (!s.empty() && (s.back() == ' ' || s.back() == '\t' || s.back() == '\r')) s.pop_back();
    return s;
}
std::string trim(const std::string& s) {
    size_t a = 0;
    while (a < s.size() && (s[a] == ' ' || s[a] == '\t')) ++a;
    return rtrim(s.substr(a));
}

// strip a '#' comment that starts the line or follows whitespace, outside quotes
std::string strip_comment(const std::string& s) {
    char q = 0;
    for (size_t i = 0; i < s.size(); ++i) {
        char c = s[i];
        if (q) {
            if (c == q) q = 0;
        } else if (c == '\'' || c == '"') {
            // only treat as a quote when it opens a token
            if (i == 0 || s[i - 1] == ' ' || s[i - 1] == '[' || s[i - 1] == '{' || s[i - 1] == ',') q = c;
        } else if (c == '#' && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '\t')) {
            return s.substr(0, i);
        }
    }
    return s;
}

NodePtr make_scalar(std::string s) {
    s = trim(s);
    auto n = std::make_shared<Node>();
    if (s.empty() || s == "~" || s == "null") return n;
    if (s.size() >= 2 && ((s.front() == '\'' && s.back() == '\'') || (s.front() == '"' && s.back() == '"')))
        s = s.substr(1, s.size() - 2);
    n->kind = Node::Scalar;
    n->scalar = s;
    return n;
}

class Parser {
   public:
    explicit Parser(const std::string& text) {
        std::istringstream in(text);
        std::string raw;
        int no = 0;
        while (std::getline(in, raw)) {
            ++no;
            raw_.push_back(raw);
            std::string s = rtrim(strip_comment(raw));
            int ind = 0;
            while (ind < (int)s.size() && s[ind] == ' ') ++ind;
            if (ind == (int)s.size()) {
                lines_.push_back({-1, "", no});  // blank
                continue;
            }
            if (s == "---" || s == "...") {
                lines_.push_back({-1, "", no});
                continue;
            }
            lines_.push_back({ind, s.substr(ind), no});
        }
    }

    NodePtr parse_document() {
        skip_blank();
        if (pos_ >= lines_.size()) return std::make_shared<Node>();
        return parse_block(lines_[pos_].indent);
    }

   private:
    std::vector<Line> lines_;
    std::vector<std::string> raw_;
    size_t pos_ = 0;

    [[noreturn]] void fail(const std::string& msg) const {
        int no = pos_ < lines_.size() ? lines_[pos_].lineno : -1;
        throw std::runtime_error("yaml: " + msg + " (line " + std::to_string(no) + ")");
    }
    void skip_blank() {
        while (pos_ < lines_.size() && lines_[pos_].indent < 0) ++pos_;
    }
    static bool is_seq_item(const std::string& t) { return t == "-" || t.rfind("- ", 0) == 0; }

    // position of the "key:" separator (": " or trailing ':') outside quotes/flow, or npos
    static size_t find_colon(const std::string& t) {
        char q = 0;
        int depth = 0;
        for (size_t i = 0; i < t.size(); ++i) {
            char c = t[i];
            if (q) {
                if (c == q) q = 0;
                continue;
            }
            if ((c == '\'' || c == '"') && i == 0) { q = c; continue; }
            if (c == '[' || c == '{') ++depth;
            else if (c == ']' || c == '}') --depth;
            else if (c == ':' && depth == 0 && (i + 1 == t.size() || t[i + 1] == ' ')) return i;
        }
        return std::string::npos;
    }

    NodePtr parse_block(int indent) {
        skip_blank();
        if (pos_ >= lines_.size()) return std::make_shared<Node>();
        if (is_seq_item(lines_[pos_].text)) return parse_seq(indent);
        return parse_map(indent);
    }

    NodePtr parse_seq(int indent) {
        auto n = std::make_shared<Node>();
        n->kind = Node::Seq;
        while (true) {
            skip_blank();
            if (pos_ >= lines_.size()) break;
            Line& L = lines_[pos_];
            if (L.indent != indent || !is_seq_item(L.text)) break;
            std::string rest = L.text.size() > 1 ? L.text.substr(2) : "";
            size_t lead = 0;
            while (lead < rest.size() && rest[lead] == ' ') ++lead;
            rest = rest.substr(lead);
            if (rest.empty()) {
                ++pos_;
                skip_blank();
                if (pos_ < lines_.size() && lines_[pos_].indent > indent)
                    n->seq.push_back(parse_block(lines_[pos_].indent));
                else
                    n->seq.push_back(std::make_shared<Node>());
            } else if (rest[0] == '[' || rest[0] == '{') {
                n->seq.push_back(parse_flow_from_line(rest));
            } else if (find_colon(rest) != std::string::npos) {
                // inline mapping start: re-interpret this line as a map line at deeper indent
                L.indent = indent + 2 + (int)lead;
                L.text = rest;
                n->seq.push_back(parse_map(L.indent));
            } else {
                ++pos_;
                n->seq.push_back(make_scalar(rest));
            }
        }
        return n;
    }

    NodePtr parse_map(int indent) {
        auto n = std::make_shared<Node>();
        n->kind = Node::Map;
        while (true) {
            skip_blank();
            if (pos_ >= lines_.size()) break;
            Line& L = lines_[pos_];
            if (L.indent != indent || is_seq_item(L.text)) break;
            size_t c = find_colon(L.text);
            if (c == std::string::npos) fail("expected 'key: value' in '" + L.text + "'");
            std::string key = trim(L.text.substr(0, c));
            if (key.size() >= 2 && (key.front() == '\'' || key.front() == '"')) key = key.substr(1, key.size() - 2);
            std::string val = trim(c + 1 < L.text.size() ? L.text.substr(c + 1) : "");
            NodePtr v;
            if (val.empty()) {
                ++pos_;
                skip_blank();
                if (pos_ < lines_.size() && lines_[pos_].indent > indent)
                    v = parse_block(lines_[pos_].indent);
                else if (pos_ < lines_.size() && lines_[pos_].indent == indent && is_seq_item(lines_[pos_].text))
                    v = parse_seq(indent);
                else
                    v = std::make_shared<Node>();
            } else if (val[0] == '[' || val[0] == '{') {
                v = parse_flow_from_line(val);
            } else if (val[0] == '|' || val[0] == '>') {
                v = parse_block_scalar(indent, val[0] == '>');
            } else {
                ++pos_;
                // plain multi-line scalar continuation (deeper-indented, non key lines)
                std::string acc = val;
                while (pos_ < lines_.size() && lines_[pos_].indent > indent &&
                       find_colon(lines_[pos_].text) == std::string::npos && !is_seq_item(lines_[pos_].text)) {
                    acc += " " + lines_[pos_].text;
                    ++pos_;
                }
                v = make_scalar(acc);
            }
            n->map.emplace_back(key, v);
        }
        return n;
    }

    NodePtr parse_block_scalar(int indent, bool folded) {
        ++pos_;
        std::string acc;
        while (pos_ < lines_.size()) {
            const Line& L = lines_[pos_];
            const std::string& raw = raw_[pos_];
            int ind = 0;
            while (ind < (int)raw.size() && raw[ind] == ' ') ++ind;
            bool blank = ind == (int)raw.size();
            if (!blank && ind <= indent) break;
            (void)L;
            if (!acc.empty()) acc += folded ? " " : "\n";
            acc += blank ? "" : trim(raw);
            ++pos_;
        }
        return make_scalar(acc);
    }

    // flow collection that starts in `first` (rest of current line) and may span lines
    NodePtr parse_flow_from_line(const std::string& first) {
        std::string buf = first;
        ++pos_;
        auto balanced = [](const std::string& s) {
            int d = 0;
            char q = 0;
            for (size_t i = 0; i < s.size(); ++i) {
                char c = s[i];
                if (q) { if (c == q) q = 0; continue; }
                if ((c == '\'' || c == '"') && (i == 0 || s[i - 1] == ' ' || s[i - 1] == '[' || s[i - 1] == '{' || s[i - 1] == ',')) { q = c; continue; }
                if (c == '[' || c == '{') ++d;
                if (c == ']' || c == '}') --d;
            }
            return d == 0;
        };
        while (!balanced(buf)) {
            if (pos_ >= lines_.size()) fail("unterminated flow collection");
            if (lines_[pos_].indent >= 0) buf += " " + lines_[pos_].text;
            ++pos_;
        }
        size_t p = 0;
        NodePtr n = parse_flow(buf, p);
        return n;
    }

    static void skip_ws(const std::string& s, size_t& p) {
        while (p < s.size() && (s[p] == ' ' || s[p] == '\t')) ++p;
    }

    NodePtr parse_flow(const std::string& s, size_t& p) {
        skip_ws(s, p);
        if (p >= s.size()) return std::make_shared<Node>();
        if (s[p] == '[') {
            auto n = std::make_shared<Node>();
            n->kind = Node::Seq;
            ++p;
            while (true) {
                skip_ws(s, p);
                if (p >= s.size()) fail("unterminated '['");
                if (s[p] == ']') { ++p; break; }
                n->seq.push_back(parse_flow(s, p));
                skip_ws(s, p);
                if (p < s.size() && s[p] == ',') ++p;
            }
            return n;
        }
        if (s[p] == '{') {
            auto n = std::make_shared<Node>();
            n->kind = Node::Map;
            ++p;
            while (true) {
                skip_ws(s, p);
                if (p >= s.size()) fail("unterminated '{'");
                if (s[p] == '}') { ++p; break; }
                // key up to ':' followed by space
                size_t k0 = p;
                while (p < s.size() && !(s[p] == ':' && (p + 1 >= s.size() || s[p + 1] == ' '))) ++p;
                if (p >= s.size()) fail("flow map key without ':'");
                std::string key = trim(s.substr(k0, p - k0));
                if (key.size() >= 2 && (key.front() == '\'' || key.front() == '"')) key = key.substr(1, key.size() - 2);
                ++p;
                NodePtr v = parse_flow(s, p);
                n->map.emplace_back(key, v);
                skip_ws(s, p);
                if (p < s.size() && s[p] == ',') ++p;
            }
            return n;
        }
        // scalar until , ] } at depth 0 (parentheses in species names are ordinary characters)
        size_t a = p;
        if (s[p] == '\'' || s[p] == '"') {
            char q = s[p++];
            while (p < s.size() && s[p] != q) ++p;
            if (p < s.size()) ++p;
            return make_scalar(s.substr(a, p - a));
        }
        while (p < s.size() && s[p] != ',' && s[p] != ']' && s[p] != '}') ++p;
        return make_scalar(s.substr(a, p - a));
    }
};

}  // namespace

NodePtr parse(const std::string& text) {
    Parser p(text);
    return p.parse_document();
}

NodePtr parse_file(const std::string& path) {
    std::ifstream f(path);
    if (!f) throw std::runtime_error("cannot open mechanism file '" + path + "'");
    std::stringstream ss;
    ss << f.rdbuf();
    return parse(ss.str());
}

}  // namespace yaml
}  // namespace dfchem
