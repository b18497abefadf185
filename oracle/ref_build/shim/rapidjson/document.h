// Build shim (test infrastructure): the subset of the RapidJSON API used by
// /root/reference/src/procedural/cube_world.cc, implemented over mini_json.h.
#ifndef ORACLE_SHIM_RAPIDJSON_DOCUMENT_H
#define ORACLE_SHIM_RAPIDJSON_DOCUMENT_H
#include <memory>
#include <istream>
#include "../../../../gpu-ray-tracer_b200/host/mini_json.h"
namespace rapidjson {
typedef unsigned SizeType;
class IStreamWrapper;
class Value {
protected:
    const mini_json::Value* v_;
    mutable std::vector<std::unique_ptr<Value>> kids_;  // keeps child wrappers alive for reference returns
    const Value& wrap(const mini_json::Value& c) const {
        kids_.emplace_back(new Value(&c));
        return *kids_.back();
    }
public:
    Value() : v_(nullptr) {}
    explicit Value(const mini_json::Value* v) : v_(v) {}
    bool HasMember(const char* k) const { return v_->has(k); }
    const Value& operator[](const char* k) const { return wrap(v_->at(std::string(k))); }
    const Value& operator[](int i) const { return wrap(v_->at((size_t)i)); }
    const Value& operator[](SizeType i) const { return wrap(v_->at((size_t)i)); }
    SizeType Size() const { return (SizeType)v_->size(); }
    int GetInt() const { return v_->as_int(); }
    float GetFloat() const { return v_->as_float(); }
    double GetDouble() const { return v_->as_double(); }
    const char* GetString() const { return v_->as_string().c_str(); }
};
class Document : public Value {
    std::shared_ptr<mini_json::Value> root_;
public:
    template <typename Stream> Document& ParseStream(Stream& s) {
        root_ = mini_json::parse_stream(s.stream());
        v_ = root_.get();
        return *this;
    }
};
}
#endif
