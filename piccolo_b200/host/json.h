// Minimal JSON reader for the asset loader (objects, arrays, strings, numbers, booleans, null).
// Numbers are parsed with strtod and kept as double; the loader casts to float / int exactly like the
// reference's Serializer::read does (engine/source/runtime/core/meta/serializer/serializer.cpp:21-52).
#pragma once
#include <map>
#include <memory>
#include <string>
#include <vector>

namespace pcljson
{
    class Value
    {
    public:
        enum Type { NUL, BOOL, NUMBER, STRING, ARRAY, OBJECT };
        Type                         type = NUL;
        bool                         b    = false;
        double                       num  = 0.0;
        std::string                  str;
        std::vector<Value>           arr;
        std::map<std::string, Value> obj;

        bool is_null() const { return type == NUL; }
        bool is_object() const { return type == OBJECT; }
        bool is_array() const { return type == ARRAY; }
        // obj["missing"] and arr[out of range] give a shared null value (like json11's operator[])
        const Value& operator[](const std::string& key) const;
        const Value& operator[](size_t i) const;
        size_t       size() const { return type == ARRAY ? arr.size() : 0; }
    };

    // returns false and fills err on malformed input
    bool parse(const std::string& text, Value& out, std::string& err);
    bool parse_file(const std::string& path, Value& out, std::string& err);
} // namespace pcljson
