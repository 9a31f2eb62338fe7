#include "json.h"

#include <cstdlib>
#include <fstream>
#include <sstream>

namespace pcljson
{
    static const Value g_null;

    const Value& Value::operator[](const std::string& key) const
    {
        if (type != OBJECT)
            return g_null;
        auto it = obj.find(key);
        return it == obj.end() ? g_null : it->second;
    }
    const Value& Value::operator[](size_t i) const { return (type == ARRAY && i < arr.size()) ? arr[i] : g_null; }

    namespace
    {
        struct Parser
        {
            const std::string& s;
            size_t             i = 0;
            std::string        err;
            explicit Parser(const std::string& t) : s(t) {}

            void ws()
            {
                while (i < s.size() && (s[i] == ' ' || s[i] == '\n' || s[i] == '\r' || s[i] == '\t'))
                    i++;
            }
            bool fail(const char* m)
            {
                if (err.empty())
                    err = std::string(m) + " at offset " + std::to_string(i);
                return false;
            }
            bool lit(const char* w)
            {
                size_t n = 0;
                while (w[n])
                    n++;
                if (s.compare(i, n, w) != 0)
                    return fail("bad literal");
                i += n;
                return true;
            }
            static void utf8(std::string& o, unsigned cp)
            {
                if (cp < 0x80)
                    o += (char)cp;
                else if (cp < 0x800)
                    o += (char)(0xC0 | (cp >> 6)), o += (char)(0x80 | (cp & 0x3F));
                else
                    o += (char)(0xE0 | (cp >> 12)), o += (char)(0x80 | ((cp >> 6) & 0x3F)), o += (char)(0x80 | (cp & 0x3F));
            }
            bool string(std::string& o)
            {
                if (s[i] != '"')
                    return fail("expected string");
                i++;
                while (i < s.size() && s[i] != '"')
                {
                    char c = s[i++];
                    if (c != '\\')
                    {
                        o += c;
                        continue;
                    }
                    if (i >= s.size())
                        return fail("bad escape");
                    char e = s[i++];
                    switch (e)
                    {
                        case 'n': o += '\n'; break;
                        case 't': o += '\t'; break;
                        case 'r': o += '\r'; break;
                        case 'b': o += '\b'; break;
                        case 'f': o += '\f'; break;
                        case 'u': {
                            if (i + 4 > s.size())
                                return fail("bad \\u escape");
                            utf8(o, (unsigned)std::strtoul(s.substr(i, 4).c_str(), nullptr, 16));
                            i += 4;
                            break;
                        }
                        default: o += e; // \" \\ \/
                    }
                }
                if (i >= s.size())
                    return fail("unterminated string");
                i++;
                return true;
            }
            bool value(Value& v, int depth)
            {
                if (depth > 200)
                    return fail("nesting too deep");
                ws();
                if (i >= s.size())
                    return fail("unexpected end");
                char c = s[i];
                if (c == '{')
                {
                    v.type = Value::OBJECT;
                    i++;
                    ws();
                    if (i < s.size() && s[i] == '}')
                    {
                        i++;
                        return true;
                    }
                    for (;;)
                    {
                        ws();
                        std::string k;
                        if (i >= s.size() || !string(k))
                            return fail("expected key");
                        ws();
                        if (i >= s.size() || s[i] != ':')
                            return fail("expected ':'");
                        i++;
                        if (!value(v.obj[k], depth + 1))
                            return false;
                        ws();
                        if (i < s.size() && s[i] == ',')
                        {
                            i++;
                            continue;
                        }
                        if (i < s.size() && s[i] == '}')
                        {
                            i++;
                            return true;
                        }
                        return fail("expected ',' or '}'");
                    }
                }
                if (c == '[')
                {
                    v.type = Value::ARRAY;
                    i++;
                    ws();
                    if (i < s.size() && s[i] == ']')
                    {
                        i++;
                        return true;
                    }
                    for (;;)
                    {
                        v.arr.emplace_back();
                        if (!value(v.arr.back(), depth + 1))
                            return false;
                        ws();
                        if (i < s.size() && s[i] == ',')
                        {
                            i++;
                            continue;
                        }
                        if (i < s.size() && s[i] == ']')
                        {
                            i++;
                            return true;
                        }
                        return fail("expected ',' or ']'");
                    }
                }
                if (c == '"')
                {
                    v.type = Value::STRING;
                    return string(v.str);
                }
                if (c == 't')
                {
                    v.type = Value::BOOL, v.b = true;
                    return lit("true");
                }
                if (c == 'f')
                {
                    v.type = Value::BOOL, v.b = false;
                    return lit("false");
                }
                if (c == 'n')
                {
                    v.type = Value::NUL;
                    return lit("null");
                }
                // number (also accepts what python's json writes for non-finite floats)
                const char* b = s.c_str() + i;
                char*       e = nullptr;
                double      d = std::strtod(b, &e);
                if (e == b)
                    return fail("unexpected character");
                v.type = Value::NUMBER, v.num = d;
                i += (size_t)(e - b);
                return true;
            }
        };
    } // namespace

    bool parse(const std::string& text, Value& out, std::string& err)
    {
        Parser p(text);
        out = Value();
        if (!p.value(out, 0))
        {
            err = p.err;
            return false;
        }
        p.ws();
        if (p.i != text.size())
        {
            err = "trailing characters at offset " + std::to_string(p.i);
            return false;
        }
        return true;
    }

    bool parse_file(const std::string& path, Value& out, std::string& err)
    {
        std::ifstream f(path);
        if (!f)
        {
            err = "open file: " + path + " failed!";
            return false;
        }
        std::stringstream ss;
        ss << f.rdbuf();
        return parse(ss.str(), out, err);
    }
} // namespace pcljson
