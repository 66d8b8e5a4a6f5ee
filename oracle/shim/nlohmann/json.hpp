// TEST INFRASTRUCTURE ONLY (oracle/): a stand-in for <nlohmann/json.hpp>, which this image does not have, so that two
// reference sources that only TOUCH json on their way to the arithmetic can be compiled where they lie:
//   tools/starkpil/bctree/build_const_tree.cpp  (reads {"nBits","nBitsExt","verificationHashType"} from the starkStruct
//                                                file, writes {"constRoot":[..]} to the verification key)
//   src/starkpil/fri/friProof.hpp               (proof -> json serialisers, never called by the oracle)
// Values: null, unsigned integer, string, array, object; parser for flat and nested objects of those; dump.
#pragma once
#include <cstdint>
#include <iomanip>
#include <istream>
#include <ostream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace nlohmann
{
class json
{
    enum Kind { Null, Num, Str, Arr, Obj } k = Null;
    uint64_t num = 0;
    std::string str;
    std::vector<json> arr;
    std::vector<std::pair<std::string, json>> obj;

    static void skip(std::istream &in)
    {
        while (in.good() && isspace(in.peek())) in.get();
    }
    static std::string parse_string(std::istream &in)
    {
        std::string s;
        if (in.get() != '"') throw std::runtime_error("json shim: expected '\"'");
        for (int c = in.get(); c != '"'; c = in.get()) {
            if (c == EOF) throw std::runtime_error("json shim: unterminated string");
            if (c == '\\') c = in.get();
            s.push_back((char)c);
        }
        return s;
    }
    void parse(std::istream &in)
    {
        skip(in);
        const int c = in.peek();
        if (c == '{') {
            in.get();
            k = Obj;
            skip(in);
            if (in.peek() == '}') { in.get(); return; }
            for (;;) {
                skip(in);
                std::string key = parse_string(in);
                skip(in);
                if (in.get() != ':') throw std::runtime_error("json shim: expected ':'");
                json v;
                v.parse(in);
                obj.emplace_back(key, v);
                skip(in);
                const int d = in.get();
                if (d == '}') return;
                if (d != ',') throw std::runtime_error("json shim: expected ',' or '}'");
            }
        } else if (c == '[') {
            in.get();
            k = Arr;
            skip(in);
            if (in.peek() == ']') { in.get(); return; }
            for (;;) {
                json v;
                v.parse(in);
                arr.push_back(v);
                skip(in);
                const int d = in.get();
                if (d == ']') return;
                if (d != ',') throw std::runtime_error("json shim: expected ',' or ']'");
            }
        } else if (c == '"') {
            k = Str;
            str = parse_string(in);
        } else if (isdigit(c)) {
            k = Num;
            in >> num;
        } else {
            std::string w;
            while (in.good() && isalpha(in.peek())) w.push_back((char)in.get());
            if (w == "true") { k = Num; num = 1; }
            else if (w == "false") { k = Num; num = 0; }
            else if (w != "null") throw std::runtime_error("json shim: unexpected token");
        }
    }
    void dump(std::ostream &out, int indent, int level) const
    {
        const std::string pad(indent * (level + 1), ' '), pad0(indent * level, ' ');
        const char *nl = indent ? "\n" : "";
        switch (k) {
        case Null: out << "null"; break;
        case Num: out << num; break;
        case Str: out << '"' << str << '"'; break;
        case Arr:
            out << '[' << nl;
            for (size_t i = 0; i < arr.size(); i++) { out << pad; arr[i].dump(out, indent, level + 1); out << (i + 1 < arr.size() ? "," : "") << nl; }
            out << pad0 << ']';
            break;
        case Obj:
            out << '{' << nl;
            for (size_t i = 0; i < obj.size(); i++) { out << pad << '"' << obj[i].first << "\": "; obj[i].second.dump(out, indent, level + 1); out << (i + 1 < obj.size() ? "," : "") << nl; }
            out << pad0 << '}';
            break;
        }
    }

public:
    json() {}
    json(uint64_t v) : k(Num), num(v) {}
    json(int v) : k(Num), num((uint64_t)v) {}
    json(unsigned v) : k(Num), num(v) {}
    json(long long v) : k(Num), num((uint64_t)v) {}
    json(unsigned long long v) : k(Num), num(v) {}
    json(const std::string &s) : k(Str), str(s) {}
    json(const char *s) : k(Str), str(s) {}
    static json array() { json j; j.k = Arr; return j; }
    static json object() { json j; j.k = Obj; return j; }
    json &operator[](const std::string &key)
    {
        if (k == Null) k = Obj;
        for (auto &p : obj) if (p.first == key) return p.second;
        obj.emplace_back(key, json());
        return obj.back().second;
    }
    json &operator[](const char *key) { return (*this)[std::string(key)]; }
    json &operator[](size_t i)
    {
        if (k == Null) k = Arr;
        if (arr.size() <= i) arr.resize(i + 1);
        return arr[i];
    }
    json &operator[](int i) { return (*this)[(size_t)i]; }
    void push_back(const json &v)
    {
        if (k == Null) k = Arr;
        arr.push_back(v);
    }
    size_t size() const { return k == Arr ? arr.size() : obj.size(); }
    void erase(size_t i) { if (k == Arr && i < arr.size()) arr.erase(arr.begin() + (long)i); }
    void erase(const char *key)
    {
        for (size_t i = 0; i < obj.size(); i++)
            if (obj[i].first == key) { obj.erase(obj.begin() + (long)i); return; }
    }
    operator uint64_t() const
    {
        if (k != Num) throw std::runtime_error("json shim: not a number");
        return num;
    }
    operator std::string() const { return str; }
    bool operator==(const char *s) const { return k == Str && str == s; }
    bool operator==(const std::string &s) const { return k == Str && str == s; }
    friend std::istream &operator>>(std::istream &in, json &j) { j = json(); j.parse(in); return in; }
    friend std::ostream &operator<<(std::ostream &out, const json &j)
    {
        const int w = (int)out.width(0);
        j.dump(out, w > 0 ? w : 0, 0);
        return out;
    }
};
using ordered_json = json;
} // namespace nlohmann
