// json.h — minimal JSON reader for engine presets.
// Replaces the reference's common/JsonParser (engine-sim-bridge/src/common/JsonParser.cpp)
// for the one thing the hot path needs: reading a fully resolved *.preset.json.
#pragma once
#include <cmath>
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace esb {

struct JsonError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

class Json {
public:
    enum Type { Null, Bool, Number, String, Array, Object };
    Type type = Null;
    double num = 0.0;
    bool boolean = false;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    bool isObject() const { return type == Object; }
    bool isArray() const { return type == Array; }
    bool isNumber() const { return type == Number; }
    size_t size() const { return type == Array ? arr.size() : obj.size(); }
    bool has(const std::string &k) const { return find(k) != nullptr; }
    const Json *find(const std::string &k) const {
        if (type != Object) return nullptr;
        for (auto &kv : obj) if (kv.first == k) return &kv.second;
        return nullptr;
    }
    const Json &at(const std::string &k, const std::string &ctx) const {
        const Json *j = find(k);
        if (!j) throw JsonError("Missing required field '" + k + "' in " + ctx);
        return *j;
    }
    const Json &operator[](size_t i) const { return arr.at(i); }
    double number(const std::string &k, const std::string &ctx) const {
        const Json &j = at(k, ctx);
        if (j.type == Bool) return j.boolean ? 1.0 : 0.0;
        if (j.type != Number) throw JsonError("Field '" + k + "' is not a number in " + ctx);
        return j.num;
    }
    double numberOr(const std::string &k, double dflt) const {
        const Json *j = find(k);
        return (j && j->type == Number) ? j->num : dflt;
    }
    std::string stringOr(const std::string &k, const std::string &dflt) const {
        const Json *j = find(k);
        return (j && j->type == String) ? j->str : dflt;
    }

    static Json parse(const std::string &text) {
        Parser p{text, 0};
        Json v = p.value();
        p.ws();
        if (p.pos != text.size()) throw JsonError("Trailing characters in JSON");
        return v;
    }

private:
    struct Parser {
        const std::string &s;
        size_t pos;
        void ws() { while (pos < s.size() && (s[pos] == ' ' || s[pos] == '\n' || s[pos] == '\t' || s[pos] == '\r')) ++pos; }
        char peek() { ws(); if (pos >= s.size()) throw JsonError("Unexpected end of JSON"); return s[pos]; }
        void expect(char c) { if (peek() != c) throw JsonError(std::string("Expected '") + c + "' at offset " + std::to_string(pos)); ++pos; }
        Json value() {
            char c = peek();
            Json v;
            if (c == '{') {
                ++pos; v.type = Object;
                if (peek() == '}') { ++pos; return v; }
                for (;;) {
                    std::string k = string();
                    expect(':');
                    v.obj.emplace_back(std::move(k), value());
                    char d = peek(); ++pos;
                    if (d == '}') break;
                    if (d != ',') throw JsonError("Expected ',' or '}' at offset " + std::to_string(pos));
                }
            } else if (c == '[') {
                ++pos; v.type = Array;
                if (peek() == ']') { ++pos; return v; }
                for (;;) {
                    v.arr.push_back(value());
                    char d = peek(); ++pos;
                    if (d == ']') break;
                    if (d != ',') throw JsonError("Expected ',' or ']' at offset " + std::to_string(pos));
                }
            } else if (c == '"') {
                v.type = String; v.str = string();
            } else if (s.compare(pos, 4, "true") == 0) { v.type = Bool; v.boolean = true; pos += 4; }
            else if (s.compare(pos, 5, "false") == 0) { v.type = Bool; v.boolean = false; pos += 5; }
            else if (s.compare(pos, 4, "null") == 0) { v.type = Null; pos += 4; }
            else {
                const char *b = s.c_str() + pos; char *e = nullptr;
                v.num = std::strtod(b, &e);
                if (e == b) throw JsonError("Invalid JSON value at offset " + std::to_string(pos));
                v.type = Number; pos += (size_t)(e - b);
            }
            return v;
        }
        std::string string() {
            expect('"');
            std::string out;
            while (pos < s.size() && s[pos] != '"') {
                char c = s[pos++];
                if (c == '\\' && pos < s.size()) {
                    char e = s[pos++];
                    switch (e) {
                    case 'n': out += '\n'; break; case 't': out += '\t'; break;
                    case 'r': out += '\r'; break; case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': pos += 4; out += '?'; break;
                    default: out += e;
                    }
                } else out += c;
            }
            if (pos >= s.size()) throw JsonError("Unterminated string in JSON");
            ++pos;
            return out;
        }
    };
};

} // namespace esb
