// json_min.hpp -- a small JSON DOM (parser + pretty printer) for the PPQSFLHE file formats.
//
// The reference reads and writes its files with nlohmann::json (utils/json, an un-vendored submodule absent from
// /root/reference): configs, the `weights_summary` containers (server/src/changeCipherDomain.cpp:48-55,118-124)
// and -- through cereal's JSON archive -- CC.json and the key files.  This header is an independent, minimal
// implementation of the subset those files need:
//   * numbers keep their source text, so 64-bit residues (up to 2^60) and doubles both survive unchanged;
//   * objects print with keys in lexicographic order and a 2-space indent, which is what
//     `out << std::setw(2) << json` produces with nlohmann's default std::map-backed object type.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace jmin {

struct Value;
typedef std::shared_ptr<Value> Ptr;

struct Value {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    std::string s;                                   // String payload or Number source text
    std::vector<Ptr> arr;
    std::vector<std::pair<std::string, Ptr>> obj;    // insertion order (cereal JSON relies on order)
    // parse_lazy only: a long array of unsigned integers (the residues of a key file: 16384 numbers per limb, 27 MB per
    // re-encryption key) is kept as its source text [raw_b, raw_e) instead of one node per number; the text must outlive
    // the tree.  numbers_to() converts it; arr stays empty.
    const char* raw_b = nullptr;
    const char* raw_e = nullptr;
    size_t raw_n = 0;

    bool is_object() const { return kind == Object; }
    bool is_array() const { return kind == Array; }
    bool is_string() const { return kind == String; }
    bool is_number() const { return kind == Number; }

    const Value* find(const std::string& key) const {
        if (kind != Object) return nullptr;
        for (const auto& kv : obj)
            if (kv.first == key) return kv.second.get();
        return nullptr;
    }
    const Value& at(const std::string& key) const {
        const Value* v = find(key);
        if (!v) throw std::runtime_error("JSON: missing key \"" + key + "\"");
        return *v;
    }
    const Value& at(size_t i) const {
        if (kind == Array && raw_b) throw std::runtime_error("JSON: element access into a lazily held numeric array");
        if (kind != Array || i >= arr.size()) throw std::runtime_error("JSON: array index out of range");
        return *arr[i];
    }
    size_t size() const { return kind == Array ? (raw_b ? raw_n : arr.size()) : kind == Object ? obj.size() : 0; }
    // every element of an array of unsigned integers, whichever way it is held
    void numbers_to(uint64_t* out) const {
        if (kind != Array) throw std::runtime_error("JSON: not an array");
        if (!raw_b) {
            for (size_t i = 0; i < arr.size(); ++i) out[i] = arr[i]->as_u64();
            return;
        }
        size_t n = 0;
        for (const char* p = raw_b; p < raw_e;) {
            const unsigned d = (unsigned char)*p - '0';
            if (d > 9) {
                ++p;
                continue;
            }
            uint64_t v = 0;
            for (; p < raw_e && (unsigned)((unsigned char)*p - '0') <= 9; ++p) v = v * 10 + ((unsigned char)*p - '0');
            out[n++] = v;
        }
        if (n != raw_n) throw std::runtime_error("JSON: numeric array changed under the tree");
    }
    uint64_t as_u64() const {
        if (kind == String || kind == Number) return std::strtoull(s.c_str(), nullptr, 10);
        throw std::runtime_error("JSON: not a number");
    }
    int64_t as_i64() const {
        if (kind == String || kind == Number) return std::strtoll(s.c_str(), nullptr, 10);
        throw std::runtime_error("JSON: not a number");
    }
    double as_double() const {
        if (kind == Number || kind == String) return std::strtod(s.c_str(), nullptr);
        throw std::runtime_error("JSON: not a number");
    }
    const std::string& as_string() const {
        if (kind != String) throw std::runtime_error("JSON: not a string");
        return s;
    }
};

inline Ptr make(Value::Kind k) {
    Ptr p(new Value);
    p->kind = k;
    return p;
}
inline Ptr make_string(const std::string& s) {
    Ptr p = make(Value::String);
    p->s = s;
    return p;
}
inline Ptr make_number_text(const std::string& s) {
    Ptr p = make(Value::Number);
    p->s = s;
    return p;
}
inline Ptr make_u64(uint64_t v) { return make_number_text(std::to_string(v)); }
inline Ptr make_double(double v) {
    char buf[40];
    std::snprintf(buf, sizeof buf, "%.17g", v);
    std::string t(buf);
    if (t.find_first_of(".eEn") == std::string::npos) t += ".0";  // keep it a float on re-read
    return make_number_text(t);
}
inline void set(Ptr& o, const std::string& key, Ptr v) {
    for (auto& kv : o->obj)
        if (kv.first == key) {
            kv.second = v;
            return;
        }
    o->obj.emplace_back(key, v);
}

class Parser {
  public:
    Parser(const char* p, size_t n, bool lazy = false) : p_(p), e_(p + n), lazy_(lazy) {}
    Ptr parse() {
        Ptr v = value();
        ws();
        if (p_ != e_) fail("trailing characters");
        return v;
    }

  private:
    const char* p_;
    const char* e_;
    bool lazy_ = false;
    [[noreturn]] void fail(const char* m) { throw std::runtime_error(std::string("JSON parse error: ") + m); }
    void ws() {
        while (p_ < e_ && (*p_ == ' ' || *p_ == '\n' || *p_ == '\t' || *p_ == '\r')) ++p_;
    }
    Ptr value() {
        ws();
        if (p_ >= e_) fail("unexpected end");
        switch (*p_) {
            case '{': return object();
            case '[': return array();
            case '"': return make_string(string());
            case 't': lit("true"); { Ptr v = make(Value::Bool); v->b = true; return v; }
            case 'f': lit("false"); { Ptr v = make(Value::Bool); v->b = false; return v; }
            case 'n': lit("null"); return make(Value::Null);
            default: return number();
        }
    }
    void lit(const char* t) {
        size_t n = std::strlen(t);
        if ((size_t)(e_ - p_) < n || std::memcmp(p_, t, n) != 0) fail("bad literal");
        p_ += n;
    }
    Ptr number() {
        const char* s = p_;
        while (p_ < e_ && (std::strchr("+-0123456789.eE", *p_) != nullptr)) ++p_;
        if (s == p_) fail("bad value");
        return make_number_text(std::string(s, p_));
    }
    std::string string() {
        ++p_;  // opening quote
        std::string out;
        // fast path: base64 payloads are megabytes without escapes
        const char* s = p_;
        {
            const char* q = static_cast<const char*>(std::memchr(p_, '"', (size_t)(e_ - p_)));
            if (!q) fail("unterminated string");
            const char* b = static_cast<const char*>(std::memchr(p_, '\\', (size_t)(q - p_)));
            p_ = b ? b : q;
        }
        out.assign(s, p_);
        while (p_ < e_ && *p_ != '"') {
            if (*p_ == '\\') {
                if (++p_ >= e_) fail("bad escape");
                switch (*p_) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': {
                        if (e_ - p_ < 5) fail("bad \\u escape");
                        unsigned cp = (unsigned)std::strtoul(std::string(p_ + 1, p_ + 5).c_str(), nullptr, 16);
                        p_ += 4;
                        if (cp < 0x80) out += (char)cp;
                        else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
                        else { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
                        break;
                    }
                    default: out += *p_;  // \" \\ \/
                }
                ++p_;
            } else {
                out += *p_++;
            }
        }
        if (p_ >= e_) fail("unterminated string");
        ++p_;
        return out;
    }
    Ptr array() {
        ++p_;
        Ptr v = make(Value::Array);
        ws();
        if (p_ < e_ && *p_ == ']') { ++p_; return v; }
        if (lazy_ && p_ < e_ && (unsigned)((unsigned char)*p_ - '0') <= 9) {
            // digits, commas and white space up to the closing bracket: an array of unsigned integers.  One branch-free
            // pass classifies the bytes and counts the commas; numbers_to() checks the element count again
            const char* s = p_;
            const char* q = static_cast<const char*>(std::memchr(s, ']', (size_t)(e_ - s)));
            if (q) {
                static const struct Cls {
                    unsigned char t[256];
                    Cls() {
                        for (int i = 0; i < 256; ++i) t[i] = 2;  // anything else
                        for (int i = '0'; i <= '9'; ++i) t[i] = 0;
                        t[(unsigned char)' '] = t[(unsigned char)'\n'] = t[(unsigned char)'\t'] = t[(unsigned char)'\r'] = 0;
                        t[(unsigned char)','] = 1;
                    }
                } cls;
                size_t commas = 0;
                unsigned bad = 0;
                for (const char* r = s; r < q; ++r) {
                    const unsigned char c = cls.t[(unsigned char)*r];
                    commas += c & 1u;
                    bad |= c & 2u;
                }
                if (!bad && commas + 1 >= 64) {
                    v->raw_b = s;
                    v->raw_e = q;
                    v->raw_n = commas + 1;
                    p_ = q + 1;
                    return v;
                }
            }
        }
        for (;;) {
            v->arr.push_back(value());
            ws();
            if (p_ >= e_) fail("unterminated array");
            if (*p_ == ',') { ++p_; continue; }
            if (*p_ == ']') { ++p_; return v; }
            fail("expected , or ]");
        }
    }
    Ptr object() {
        ++p_;
        Ptr v = make(Value::Object);
        ws();
        if (p_ < e_ && *p_ == '}') { ++p_; return v; }
        for (;;) {
            ws();
            if (p_ >= e_ || *p_ != '"') fail("expected object key");
            std::string k = string();
            ws();
            if (p_ >= e_ || *p_ != ':') fail("expected :");
            ++p_;
            v->obj.emplace_back(std::move(k), value());
            ws();
            if (p_ >= e_) fail("unterminated object");
            if (*p_ == ',') { ++p_; continue; }
            if (*p_ == '}') { ++p_; return v; }
            fail("expected , or }");
        }
    }
};

inline Ptr parse(const std::string& text) { return Parser(text.data(), text.size()).parse(); }
// the same with long arrays of unsigned integers held as text spans (Value::raw_b): `text` must outlive the tree
inline Ptr parse_lazy(const std::string& text) { return Parser(text.data(), text.size(), true).parse(); }

inline void escape_to(std::string& out, const std::string& s) {
    out += '"';
    {   // fast path: nothing to escape (base64 payloads are megabytes of plain characters)
        // eight bytes per step: a byte is < 0x20, '"' or '\\' iff the matching "has zero byte" test fires
        const uint64_t ones = 0x0101010101010101ull, highs = 0x8080808080808080ull;
        auto has_zero = [&](uint64_t v) { return ((v - ones) & ~v & highs) != 0; };
        bool plain = true;
        size_t i = 0;
        for (; plain && i + 8 <= s.size(); i += 8) {
            uint64_t w;
            std::memcpy(&w, s.data() + i, 8);
            plain = !(has_zero(w & 0xe0e0e0e0e0e0e0e0ull) || has_zero(w ^ (ones * 0x22)) || has_zero(w ^ (ones * 0x5c)));
        }
        for (; plain && i < s.size(); ++i) {
            const unsigned char c = (unsigned char)s[i];
            plain = c >= 0x20 && c != '"' && c != '\\';
        }
        if (plain) {
            out += s;
            out += '"';
            return;
        }
    }
    for (unsigned char c : s) {
        switch (c) {
            case '"': out += "\\\""; break;
            case '\\': out += "\\\\"; break;
            case '\n': out += "\\n"; break;
            case '\t': out += "\\t"; break;
            case '\r': out += "\\r"; break;
            default:
                if (c < 0x20) { char b[8]; std::snprintf(b, sizeof b, "\\u%04x", c); out += b; }
                else out += (char)c;
        }
    }
    out += '"';
}

// Pretty printer matching `os << std::setw(indent) << nlohmann::json`: sorted keys, one element per line.
inline void dump_to(std::string& out, const Value& v, int indent, int cur, bool sort_keys) {
    switch (v.kind) {
        case Value::Null: out += "null"; break;
        case Value::Bool: out += v.b ? "true" : "false"; break;
        case Value::Number: out += v.s; break;
        case Value::String: escape_to(out, v.s); break;
        case Value::Array: {
            if (v.raw_b) {  // lazily held numbers: their source text, whatever its layout
                out += '[';
                out.append(v.raw_b, v.raw_e);
                out += ']';
                break;
            }
            if (v.arr.empty()) { out += "[]"; break; }
            out += "[\n";
            for (size_t i = 0; i < v.arr.size(); ++i) {
                out.append((size_t)(cur + indent), ' ');
                dump_to(out, *v.arr[i], indent, cur + indent, sort_keys);
                if (i + 1 < v.arr.size()) out += ',';
                out += '\n';
            }
            out.append((size_t)cur, ' ');
            out += ']';
            break;
        }
        case Value::Object: {
            if (v.obj.empty()) { out += "{}"; break; }
            std::vector<const std::pair<std::string, Ptr>*> items;
            for (const auto& kv : v.obj) items.push_back(&kv);
            if (sort_keys)
                for (size_t i = 1; i < items.size(); ++i)  // insertion sort: objects here are tiny
                    for (size_t j = i; j > 0 && items[j]->first < items[j - 1]->first; --j) std::swap(items[j], items[j - 1]);
            out += "{\n";
            for (size_t i = 0; i < items.size(); ++i) {
                out.append((size_t)(cur + indent), ' ');
                escape_to(out, items[i]->first);
                out += ": ";
                dump_to(out, *items[i]->second, indent, cur + indent, sort_keys);
                if (i + 1 < items.size()) out += ',';
                out += '\n';
            }
            out.append((size_t)cur, ' ');
            out += '}';
            break;
        }
    }
}
inline std::string dump(const Value& v, int indent = 2, bool sort_keys = true) {
    std::string out;
    dump_to(out, v, indent, 0, sort_keys);
    return out;
}

}  // namespace jmin
