// Json.h -- a small JSON document model for the scene loader (host/SceneLoader.cpp).
//
// The reference parses its scene files with rapidjson using kParseCommentsFlag | kParseTrailingCommasFlag |
// kParseNanAndInfFlag (src/runtime/loader/Parser.cpp:546) and walks objects in MEMBER ORDER, repeated keys included
// (a transform object may hold two "qrotate" members, Parser.cpp:124-243).  This reader accepts the same dialect and keeps
// objects as ordered member lists for that reason.  No third-party code.
#pragma once

#include <cmath>
#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace IG {
namespace json {

struct Value;
using Member = std::pair<std::string, Value>;

struct Value {
    enum Type { Null, Bool, Number, String, Array, Object } type = Null;
    bool                boolean = false;
    double              number  = 0.0;
    std::string         string;
    std::vector<Value>  array;
    std::vector<Member> members; // in file order, repeated keys kept

    bool isNull() const { return type == Null; }
    bool isNumber() const { return type == Number; }
    bool isString() const { return type == String; }
    bool isArray() const { return type == Array; }
    bool isObject() const { return type == Object; }
    bool isBool() const { return type == Bool; }

    // first member of that name (what a name-keyed lookup sees), nullptr when absent
    const Value* find(const std::string& key) const
    {
        if (type != Object) return nullptr;
        for (const Member& m : members)
            if (m.first == key) return &m.second;
        return nullptr;
    }
    bool has(const std::string& key) const { return find(key) != nullptr; }

    double num(const std::string& key, double def) const
    {
        const Value* v = find(key);
        return v && v->isNumber() ? v->number : def;
    }
    bool flag(const std::string& key, bool def) const
    {
        const Value* v = find(key);
        if (!v) return def;
        if (v->isBool()) return v->boolean;
        if (v->isNumber()) return v->number != 0.0;
        return def;
    }
    std::string str(const std::string& key, const std::string& def) const
    {
        const Value* v = find(key);
        return v && v->isString() ? v->string : def;
    }
};

class Parser {
public:
    explicit Parser(const std::string& text) : s(text) {}

    Value parseDocument()
    {
        Value v = parseValue();
        skip();
        if (pos != s.size()) fail("trailing characters after the document");
        return v;
    }

private:
    const std::string& s;
    size_t             pos = 0;

    [[noreturn]] void fail(const std::string& what) const
    {
        size_t line = 1;
        for (size_t i = 0; i < pos && i < s.size(); ++i)
            if (s[i] == '\n') ++line;
        throw std::runtime_error("JSON: " + what + " (line " + std::to_string(line) + ")");
    }

    void skip() // white space and both comment forms
    {
        for (;;) {
            while (pos < s.size() && (s[pos] == ' ' || s[pos] == '\t' || s[pos] == '\n' || s[pos] == '\r')) ++pos;
            if (pos + 1 < s.size() && s[pos] == '/' && s[pos + 1] == '/') {
                while (pos < s.size() && s[pos] != '\n') ++pos;
            } else if (pos + 1 < s.size() && s[pos] == '/' && s[pos + 1] == '*') {
                const size_t end = s.find("*/", pos + 2);
                pos              = end == std::string::npos ? s.size() : end + 2;
            } else {
                return;
            }
        }
    }

    bool literal(const char* word)
    {
        size_t n = 0;
        while (word[n]) ++n;
        if (s.compare(pos, n, word) == 0) {
            pos += n;
            return true;
        }
        return false;
    }

    Value parseValue()
    {
        skip();
        if (pos >= s.size()) fail("unexpected end of input");
        const char c = s[pos];
        Value      v;
        if (c == '{') {
            ++pos;
            v.type = Value::Object;
            for (;;) {
                skip();
                if (pos < s.size() && s[pos] == '}') { ++pos; break; }
                if (pos >= s.size() || s[pos] != '"') fail("expected a member name");
                std::string key = parseString();
                skip();
                if (pos >= s.size() || s[pos] != ':') fail("expected ':'");
                ++pos;
                Value member = parseValue();
                v.members.emplace_back(std::move(key), std::move(member));
                skip();
                if (pos < s.size() && s[pos] == ',') { ++pos; continue; } // a trailing comma falls through to '}'
                if (pos < s.size() && s[pos] == '}') { ++pos; break; }
                fail("expected ',' or '}'");
            }
        } else if (c == '[') {
            ++pos;
            v.type = Value::Array;
            for (;;) {
                skip();
                if (pos < s.size() && s[pos] == ']') { ++pos; break; }
                v.array.push_back(parseValue());
                skip();
                if (pos < s.size() && s[pos] == ',') { ++pos; continue; }
                if (pos < s.size() && s[pos] == ']') { ++pos; break; }
                fail("expected ',' or ']'");
            }
        } else if (c == '"') {
            v.type   = Value::String;
            v.string = parseString();
        } else if (literal("true")) {
            v.type    = Value::Bool;
            v.boolean = true;
        } else if (literal("false")) {
            v.type = Value::Bool;
        } else if (literal("null")) {
        } else if (literal("NaN")) {
            v.type   = Value::Number;
            v.number = std::nan("");
        } else if (literal("Infinity") || literal("Inf")) {
            v.type   = Value::Number;
            v.number = INFINITY;
        } else if (literal("-Infinity") || literal("-Inf")) {
            v.type   = Value::Number;
            v.number = -INFINITY;
        } else {
            const char* begin = s.c_str() + pos;
            char*       end   = nullptr;
            v.number          = std::strtod(begin, &end);
            if (end == begin) fail("unexpected character");
            v.type = Value::Number;
            pos += (size_t)(end - begin);
        }
        return v;
    }

    std::string parseString()
    {
        std::string out;
        ++pos; // opening quote
        while (pos < s.size() && s[pos] != '"') {
            char c = s[pos++];
            if (c == '\\' && pos < s.size()) {
                const char e = s[pos++];
                switch (e) {
                case 'n': out += '\n'; break;
                case 't': out += '\t'; break;
                case 'r': out += '\r'; break;
                case 'b': out += '\b'; break;
                case 'f': out += '\f'; break;
                case 'u': {
                    unsigned cp = 0;
                    for (int i = 0; i < 4 && pos < s.size(); ++i) cp = cp * 16 + (unsigned)std::strtoul(std::string(1, s[pos++]).c_str(), nullptr, 16);
                    if (cp < 0x80) {
                        out += (char)cp;
                    } else if (cp < 0x800) {
                        out += (char)(0xC0 | (cp >> 6));
                        out += (char)(0x80 | (cp & 0x3F));
                    } else {
                        out += (char)(0xE0 | (cp >> 12));
                        out += (char)(0x80 | ((cp >> 6) & 0x3F));
                        out += (char)(0x80 | (cp & 0x3F));
                    }
                    break;
                }
                default: out += e; break; // \" \\ \/
                }
            } else {
                out += c;
            }
        }
        if (pos >= s.size()) fail("unterminated string");
        ++pos; // closing quote
        return out;
    }
};

inline Value parse(const std::string& text) { return Parser(text).parseDocument(); }

} // namespace json
} // namespace IG
