// mini_json.h — a small, dependency-free JSON reader for world*.json scene files.
//
// The reference parses its scene files with RapidJSON (un-vendored dependency,
// /root/reference/CMakeLists.txt:20; uses in src/procedural/cube_world.cc:1-2,23-30,38-200).
// This reader exposes the handful of accessors that loader needs. Numbers are parsed
// with strtod (correctly rounded) and kept as double; integer-valued literals also
// remember their int64 value so GetInt() is exact.
#ifndef RTB200_MINI_JSON_H
#define RTB200_MINI_JSON_H

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <istream>
#include <iterator>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace mini_json {

class Value {
public:
    enum Kind { Null, Bool, Number, String, Array, Object };

    Value() : kind_(Null), num_(0.0), inum_(0), is_int_(false), b_(false) {}

    Kind kind() const { return kind_; }
    bool is_object() const { return kind_ == Object; }
    bool is_array() const { return kind_ == Array; }
    bool is_number() const { return kind_ == Number; }

    bool has(const std::string& key) const {
        return kind_ == Object && obj_.find(key) != obj_.end();
    }
    const Value& at(const std::string& key) const {
        if (kind_ != Object) throw std::runtime_error("json: not an object (key '" + key + "')");
        auto it = obj_.find(key);
        if (it == obj_.end()) throw std::runtime_error("json: missing key '" + key + "'");
        return *it->second;
    }
    const Value& at(size_t idx) const {
        if (kind_ != Array) throw std::runtime_error("json: not an array");
        if (idx >= arr_.size()) throw std::runtime_error("json: index out of range");
        return *arr_[idx];
    }
    size_t size() const {
        if (kind_ == Array) return arr_.size();
        if (kind_ == Object) return obj_.size();
        return 0;
    }
    double as_double() const {
        if (kind_ != Number) throw std::runtime_error("json: not a number");
        return num_;
    }
    float as_float() const { return (float)as_double(); }
    int as_int() const {
        if (kind_ != Number) throw std::runtime_error("json: not a number");
        return is_int_ ? (int)inum_ : (int)num_;
    }
    bool as_bool() const { return kind_ == Bool ? b_ : false; }
    const std::string& as_string() const {
        if (kind_ != String) throw std::runtime_error("json: not a string");
        return str_;
    }
    std::vector<std::string> keys() const {
        std::vector<std::string> ks;
        for (auto& kv : obj_) ks.push_back(kv.first);
        return ks;
    }

private:
    friend class Parser;
    Kind kind_;
    double num_;
    int64_t inum_;
    bool is_int_;
    bool b_;
    std::string str_;
    std::vector<std::shared_ptr<Value>> arr_;
    std::map<std::string, std::shared_ptr<Value>> obj_;
};

class Parser {
public:
    explicit Parser(const std::string& text) : s_(text), i_(0) {}

    std::shared_ptr<Value> parse() {
        std::shared_ptr<Value> v = parse_value();
        skip_ws();
        if (i_ != s_.size()) fail("trailing characters");
        return v;
    }

private:
    const std::string& s_;
    size_t i_;

    [[noreturn]] void fail(const char* what) const {
        throw std::runtime_error(std::string("json parse error at byte ") + std::to_string(i_) + ": " + what);
    }
    void skip_ws() {
        while (i_ < s_.size() && (s_[i_] == ' ' || s_[i_] == '\t' || s_[i_] == '\n' || s_[i_] == '\r')) i_++;
    }
    char peek() {
        skip_ws();
        if (i_ >= s_.size()) fail("unexpected end of input");
        return s_[i_];
    }
    void expect(char c) {
        if (peek() != c) fail("unexpected character");
        i_++;
    }
    std::shared_ptr<Value> parse_value() {
        char c = peek();
        if (c == '{') return parse_object();
        if (c == '[') return parse_array();
        if (c == '"') {
            auto v = std::make_shared<Value>();
            v->kind_ = Value::String;
            v->str_ = parse_string();
            return v;
        }
        if (c == 't' || c == 'f' || c == 'n') return parse_literal();
        return parse_number();
    }
    std::shared_ptr<Value> parse_literal() {
        auto v = std::make_shared<Value>();
        if (s_.compare(i_, 4, "true") == 0) { v->kind_ = Value::Bool; v->b_ = true; i_ += 4; }
        else if (s_.compare(i_, 5, "false") == 0) { v->kind_ = Value::Bool; v->b_ = false; i_ += 5; }
        else if (s_.compare(i_, 4, "null") == 0) { v->kind_ = Value::Null; i_ += 4; }
        else fail("bad literal");
        return v;
    }
    std::shared_ptr<Value> parse_number() {
        size_t start = i_;
        bool is_int = true;
        if (i_ < s_.size() && (s_[i_] == '-' || s_[i_] == '+')) i_++;
        while (i_ < s_.size()) {
            char c = s_[i_];
            if (c >= '0' && c <= '9') { i_++; }
            else if (c == '.' || c == 'e' || c == 'E' || c == '-' || c == '+') { is_int = false; i_++; }
            else break;
        }
        if (i_ == start) fail("bad number");
        std::string tok = s_.substr(start, i_ - start);
        auto v = std::make_shared<Value>();
        v->kind_ = Value::Number;
        char* end = nullptr;
        v->num_ = std::strtod(tok.c_str(), &end);
        if (end == tok.c_str()) fail("bad number");
        v->is_int_ = is_int;
        if (is_int) v->inum_ = std::strtoll(tok.c_str(), nullptr, 10);
        return v;
    }
    std::string parse_string() {
        expect('"');
        std::string out;
        while (true) {
            if (i_ >= s_.size()) fail("unterminated string");
            char c = s_[i_++];
            if (c == '"') break;
            if (c == '\\') {
                if (i_ >= s_.size()) fail("bad escape");
                char e = s_[i_++];
                switch (e) {
                    case 'n': out.push_back('\n'); break;
                    case 't': out.push_back('\t'); break;
                    case 'r': out.push_back('\r'); break;
                    case 'b': out.push_back('\b'); break;
                    case 'f': out.push_back('\f'); break;
                    case 'u': {
                        if (i_ + 4 > s_.size()) fail("bad \\u escape");
                        unsigned cp = (unsigned)std::strtoul(s_.substr(i_, 4).c_str(), nullptr, 16);
                        i_ += 4;
                        if (cp < 0x80) out.push_back((char)cp);
                        else if (cp < 0x800) { out.push_back((char)(0xC0 | (cp >> 6))); out.push_back((char)(0x80 | (cp & 0x3F))); }
                        else { out.push_back((char)(0xE0 | (cp >> 12))); out.push_back((char)(0x80 | ((cp >> 6) & 0x3F))); out.push_back((char)(0x80 | (cp & 0x3F))); }
                        break;
                    }
                    default: out.push_back(e); break;
                }
            } else {
                out.push_back(c);
            }
        }
        return out;
    }
    std::shared_ptr<Value> parse_array() {
        expect('[');
        auto v = std::make_shared<Value>();
        v->kind_ = Value::Array;
        if (peek() == ']') { i_++; return v; }
        while (true) {
            v->arr_.push_back(parse_value());
            char c = peek();
            i_++;
            if (c == ']') break;
            if (c != ',') fail("expected ',' or ']'");
        }
        return v;
    }
    std::shared_ptr<Value> parse_object() {
        expect('{');
        auto v = std::make_shared<Value>();
        v->kind_ = Value::Object;
        if (peek() == '}') { i_++; return v; }
        while (true) {
            if (peek() != '"') fail("expected string key");
            std::string key = parse_string();
            expect(':');
            v->obj_[key] = parse_value();
            char c = peek();
            i_++;
            if (c == '}') break;
            if (c != ',') fail("expected ',' or '}'");
        }
        return v;
    }
};

inline std::shared_ptr<Value> parse_text(const std::string& text) {
    Parser p(text);
    return p.parse();
}

inline std::shared_ptr<Value> parse_stream(std::istream& is) {
    std::string text((std::istreambuf_iterator<char>(is)), std::istreambuf_iterator<char>());
    return parse_text(text);
}

}  // namespace mini_json

#endif
