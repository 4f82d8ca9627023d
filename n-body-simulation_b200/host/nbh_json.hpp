// nbh_json.hpp — a small JSON reader for config.json (the reference fetches nlohmann/json from the network at build
// time, CMakeLists.txt:102-107; there is no network here). Reads ONE value and stops, so the trailing '#' comment lines
// of the shipped config.json (config.json:67-68) are tolerated exactly as `in >> j` tolerates them in the reference.
#pragma once
#include <cctype>
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace nbh {

class Json {
public:
    enum Type { Null, Bool, Number, String, Array, Object };
    Type type = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Json> arr;
    std::map<std::string, Json> obj;

    bool contains(const std::string& k) const { return type == Object && obj.count(k) > 0; }
    const Json& at(const std::string& k) const {
        if (type != Object) throw std::runtime_error("config: expected an object while looking up '" + k + "'");
        auto it = obj.find(k);
        if (it == obj.end()) throw std::runtime_error("config: missing key '" + k + "'");
        return it->second;
    }
    const Json& operator[](const std::string& k) const { return at(k); }
    double as_double() const { if (type != Number) throw std::runtime_error("config: expected a number"); return num; }
    int as_int() const { return (int)as_double(); }
    long as_long() const { return (long)as_double(); }
    bool as_bool() const { if (type != Bool) throw std::runtime_error("config: expected true/false"); return b; }
    const std::string& as_string() const { if (type != String) throw std::runtime_error("config: expected a string"); return str; }
    double get(const std::string& k, double dflt) const { return contains(k) ? at(k).as_double() : dflt; }
    std::string get(const std::string& k, const std::string& dflt) const { return contains(k) ? at(k).as_string() : dflt; }

    static Json parse(const std::string& text) { size_t p = 0; Json v = value(text, p); return v; }

private:
    static void ws(const std::string& s, size_t& p) { while (p < s.size() && std::isspace((unsigned char)s[p])) ++p; }
    [[noreturn]] static void bad(const std::string& what, size_t p) { throw std::runtime_error("config: JSON syntax error (" + what + ") at offset " + std::to_string(p)); }
    static Json value(const std::string& s, size_t& p) {
        ws(s, p);
        if (p >= s.size()) bad("unexpected end", p);
        Json v;
        const char c = s[p];
        if (c == '{') {
            v.type = Object; ++p; ws(s, p);
            if (p < s.size() && s[p] == '}') { ++p; return v; }
            for (;;) {
                ws(s, p);
                if (p >= s.size() || s[p] != '"') bad("expected a key", p);
                std::string k = string(s, p);
                ws(s, p);
                if (p >= s.size() || s[p] != ':') bad("expected ':'", p);
                ++p;
                v.obj[k] = value(s, p);
                ws(s, p);
                if (p < s.size() && s[p] == ',') { ++p; continue; }
                if (p < s.size() && s[p] == '}') { ++p; return v; }
                bad("expected ',' or '}'", p);
            }
        }
        if (c == '[') {
            v.type = Array; ++p; ws(s, p);
            if (p < s.size() && s[p] == ']') { ++p; return v; }
            for (;;) {
                v.arr.push_back(value(s, p));
                ws(s, p);
                if (p < s.size() && s[p] == ',') { ++p; continue; }
                if (p < s.size() && s[p] == ']') { ++p; return v; }
                bad("expected ',' or ']'", p);
            }
        }
        if (c == '"') { v.type = String; v.str = string(s, p); return v; }
        if (s.compare(p, 4, "true") == 0) { v.type = Bool; v.b = true; p += 4; return v; }
        if (s.compare(p, 5, "false") == 0) { v.type = Bool; v.b = false; p += 5; return v; }
        if (s.compare(p, 4, "null") == 0) { p += 4; return v; }
        char* end = nullptr;
        v.num = std::strtod(s.c_str() + p, &end);
        if (end == s.c_str() + p) bad("unexpected character", p);
        v.type = Number; p = (size_t)(end - s.c_str());
        return v;
    }
    static std::string string(const std::string& s, size_t& p) {
        std::string out; ++p;
        while (p < s.size() && s[p] != '"') {
            if (s[p] == '\\' && p + 1 < s.size()) {
                const char e = s[++p];
                switch (e) { case 'n': out += '\n'; break; case 't': out += '\t'; break; case 'r': out += '\r'; break;
                             case 'b': out += '\b'; break; case 'f': out += '\f'; break; default: out += e; }
                ++p;
            } else out += s[p++];
        }
        if (p >= s.size()) bad("unterminated string", p);
        ++p;
        return out;
    }
};

} // namespace nbh
