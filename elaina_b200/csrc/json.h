// Minimal JSON document model for the configuration surface (elaina-exec conf.json, tcnn network block, colour
// files). The reference uses nlohmann/json (ext/json); this is a small independent recursive-descent parser that
// covers what those files contain: objects, arrays, strings with escapes, numbers, booleans, null.
// Lookup helpers mirror json_get_or_throw / json_get_optional (core/common.h:152-213): a missing key is an error
// the caller turns into WOST_ERR_CONFIG.
#pragma once
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace wjson
{
    struct Error : std::runtime_error
    {
        using std::runtime_error::runtime_error;
    };

    struct Value
    {
        enum Type
        {
            Null,
            Bool,
            Number,
            String,
            Array,
            Object
        };
        Type type = Null;
        bool b = false;
        double num = 0.0;
        std::string str;
        std::vector<Value> arr;
        std::vector<std::pair<std::string, Value>> obj; // insertion order kept for dump()

        bool is_object() const { return type == Object; }
        bool is_array() const { return type == Array; }
        bool is_string() const { return type == String; }
        bool is_number() const { return type == Number; }
        const Value *find(const std::string &key) const
        {
            if (type != Object)
                return nullptr;
            for (const auto &kv : obj)
                if (kv.first == key)
                    return &kv.second;
            return nullptr;
        }
        bool contains(const std::string &key) const { return find(key) != nullptr; }
        const Value &at(const std::string &key) const
        {
            const Value *v = find(key);
            if (!v)
                throw Error("Key not found: " + key);
            return *v;
        }
        const Value &at(size_t i) const
        {
            if (type != Array || i >= arr.size())
                throw Error("array index out of range");
            return arr[i];
        }
        size_t size() const { return type == Array ? arr.size() : (type == Object ? obj.size() : 0); }
        double as_number() const
        {
            if (type == Number)
                return num;
            if (type == Bool)
                return b ? 1.0 : 0.0;
            throw Error("expected a number");
        }
        int as_int() const { return (int)std::llround(as_number()); }
        float as_float() const { return (float)as_number(); }
        bool as_bool() const
        {
            if (type == Bool)
                return b;
            if (type == Number)
                return num != 0.0;
            throw Error("expected a boolean");
        }
        const std::string &as_string() const
        {
            if (type != String)
                throw Error("expected a string");
            return str;
        }
        double number_or(const std::string &key, double dflt) const
        {
            const Value *v = find(key);
            return (v && (v->type == Number || v->type == Bool)) ? v->as_number() : dflt;
        }
        std::string string_or(const std::string &key, const std::string &dflt) const
        {
            const Value *v = find(key);
            return (v && v->type == String) ? v->str : dflt;
        }
    };

    class Parser
    {
    public:
        explicit Parser(const std::string &text) : s(text) {}
        Value parse()
        {
            Value v = value();
            ws();
            if (i != s.size())
                fail("trailing characters");
            return v;
        }

    private:
        const std::string &s;
        size_t i = 0;
        [[noreturn]] void fail(const char *what) const
        {
            char buf[128];
            snprintf(buf, sizeof(buf), "JSON parse error at offset %zu: %s", i, what);
            throw Error(buf);
        }
        void ws()
        {
            while (i < s.size() && (s[i] == ' ' || s[i] == '\t' || s[i] == '\n' || s[i] == '\r'))
                ++i;
        }
        bool lit(const char *w)
        {
            size_t n = 0;
            while (w[n])
                ++n;
            if (s.compare(i, n, w) == 0)
            {
                i += n;
                return true;
            }
            return false;
        }
        Value value()
        {
            ws();
            if (i >= s.size())
                fail("unexpected end");
            Value v;
            const char c = s[i];
            if (c == '{')
            {
                ++i;
                v.type = Value::Object;
                ws();
                if (i < s.size() && s[i] == '}')
                {
                    ++i;
                    return v;
                }
                for (;;)
                {
                    ws();
                    if (i >= s.size() || s[i] != '"')
                        fail("expected object key");
                    std::string k = string();
                    ws();
                    if (i >= s.size() || s[i] != ':')
                        fail("expected ':'");
                    ++i;
                    v.obj.emplace_back(std::move(k), value());
                    ws();
                    if (i < s.size() && s[i] == ',')
                    {
                        ++i;
                        continue;
                    }
                    if (i < s.size() && s[i] == '}')
                    {
                        ++i;
                        return v;
                    }
                    fail("expected ',' or '}'");
                }
            }
            if (c == '[')
            {
                ++i;
                v.type = Value::Array;
                ws();
                if (i < s.size() && s[i] == ']')
                {
                    ++i;
                    return v;
                }
                for (;;)
                {
                    v.arr.push_back(value());
                    ws();
                    if (i < s.size() && s[i] == ',')
                    {
                        ++i;
                        continue;
                    }
                    if (i < s.size() && s[i] == ']')
                    {
                        ++i;
                        return v;
                    }
                    fail("expected ',' or ']'");
                }
            }
            if (c == '"')
            {
                v.type = Value::String;
                v.str = string();
                return v;
            }
            if (lit("true"))
            {
                v.type = Value::Bool;
                v.b = true;
                return v;
            }
            if (lit("false"))
            {
                v.type = Value::Bool;
                return v;
            }
            if (lit("null"))
                return v;
            // number
            const char *start = s.c_str() + i;
            char *end = nullptr;
            const double d = strtod(start, &end);
            if (end == start)
                fail("unexpected token");
            i += (size_t)(end - start);
            v.type = Value::Number;
            v.num = d;
            return v;
        }
        static void utf8(std::string &out, unsigned cp)
        {
            if (cp < 0x80)
                out += (char)cp;
            else if (cp < 0x800)
            {
                out += (char)(0xC0 | (cp >> 6));
                out += (char)(0x80 | (cp & 0x3F));
            }
            else
            {
                out += (char)(0xE0 | (cp >> 12));
                out += (char)(0x80 | ((cp >> 6) & 0x3F));
                out += (char)(0x80 | (cp & 0x3F));
            }
        }
        std::string string()
        {
            ++i; // opening quote
            std::string out;
            while (i < s.size() && s[i] != '"')
            {
                char c = s[i++];
                if (c == '\\')
                {
                    if (i >= s.size())
                        fail("bad escape");
                    c = s[i++];
                    switch (c)
                    {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u':
                    {
                        if (i + 4 > s.size())
                            fail("bad \\u escape");
                        utf8(out, (unsigned)strtoul(s.substr(i, 4).c_str(), nullptr, 16));
                        i += 4;
                        break;
                    }
                    default: out += c; break; // \" \\ \/
                    }
                }
                else
                    out += c;
            }
            if (i >= s.size())
                fail("unterminated string");
            ++i;
            return out;
        }
    };

    inline Value parse(const std::string &text) { return Parser(text).parse(); }

    inline void dump_string(const std::string &s, std::string &out)
    {
        out += '"';
        for (char c : s)
        {
            switch (c)
            {
            case '"': out += "\\\""; break;
            case '\\': out += "\\\\"; break;
            case '\n': out += "\\n"; break;
            case '\t': out += "\\t"; break;
            case '\r': out += "\\r"; break;
            default: out += c;
            }
        }
        out += '"';
    }
    inline void dump(const Value &v, std::string &out, int indent = 4, int level = 0)
    {
        const std::string pad((size_t)indent * (level + 1), ' '), pad0((size_t)indent * level, ' ');
        char buf[64];
        switch (v.type)
        {
        case Value::Null: out += "null"; break;
        case Value::Bool: out += v.b ? "true" : "false"; break;
        case Value::Number:
            if (v.num == std::floor(v.num) && std::fabs(v.num) < 1e15)
                snprintf(buf, sizeof(buf), "%lld", (long long)v.num);
            else
                snprintf(buf, sizeof(buf), "%.17g", v.num);
            out += buf;
            break;
        case Value::String: dump_string(v.str, out); break;
        case Value::Array:
            if (v.arr.empty())
            {
                out += "[]";
                break;
            }
            out += "[\n";
            for (size_t k = 0; k < v.arr.size(); ++k)
            {
                out += pad;
                dump(v.arr[k], out, indent, level + 1);
                out += (k + 1 < v.arr.size()) ? ",\n" : "\n";
            }
            out += pad0 + "]";
            break;
        case Value::Object:
            if (v.obj.empty())
            {
                out += "{}";
                break;
            }
            out += "{\n";
            for (size_t k = 0; k < v.obj.size(); ++k)
            {
                out += pad;
                dump_string(v.obj[k].first, out);
                out += ": ";
                dump(v.obj[k].second, out, indent, level + 1);
                out += (k + 1 < v.obj.size()) ? ",\n" : "\n";
            }
            out += pad0 + "}";
            break;
        }
    }
    inline Value make_number(double d)
    {
        Value v;
        v.type = Value::Number;
        v.num = d;
        return v;
    }
    inline Value make_string(const std::string &s)
    {
        Value v;
        v.type = Value::String;
        v.str = s;
        return v;
    }
} // namespace wjson
