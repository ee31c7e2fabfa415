// Minimal JSON reader for the reference's nnet_conf/*.json files (misc/utils.py:13-41 loads them with
// json.load).  Supports objects, arrays, strings (with escapes), numbers, true/false/null.
#pragma once
#include <stdlib.h>

#include <string>
#include <utility>
#include <vector>

namespace xv {

struct JValue {
  enum Type { Null, Bool, Num, Str, Arr, Obj } type = Null;
  bool b = false;
  double num = 0.0;
  std::string str;
  std::vector<JValue> arr;
  std::vector<std::pair<std::string, JValue>> obj;

  const JValue* get(const std::string& key) const {
    if (type != Obj) return nullptr;
    for (const auto& kv : obj)
      if (kv.first == key) return &kv.second;
    return nullptr;
  }
  bool has(const std::string& key) const { return get(key) != nullptr; }
  bool truthy() const {
    switch (type) {
      case Bool: return b;
      case Num: return num != 0.0;
      case Str: return !str.empty();
      case Arr: return !arr.empty();
      case Obj: return !obj.empty();
      default: return false;
    }
  }
};

class JParser {
 public:
  explicit JParser(const char* s) : p_(s) {}
  bool parse(JValue* out, std::string* err) {
    if (!value(out)) {
      *err = err_.empty() ? "malformed JSON" : err_;
      return false;
    }
    ws();
    if (*p_ != '\0') {
      *err = "trailing characters after JSON value";
      return false;
    }
    return true;
  }

 private:
  const char* p_;
  std::string err_;
  void ws() {
    while (*p_ == ' ' || *p_ == '\t' || *p_ == '\n' || *p_ == '\r') ++p_;
  }
  bool fail(const char* m) {
    if (err_.empty()) err_ = m;
    return false;
  }
  bool string(std::string* s) {
    if (*p_ != '"') return fail("expected string");
    ++p_;
    s->clear();
    while (*p_ && *p_ != '"') {
      if (*p_ == '\\') {
        ++p_;
        switch (*p_) {
          case 'n': s->push_back('\n'); break;
          case 't': s->push_back('\t'); break;
          case 'r': s->push_back('\r'); break;
          case 'b': s->push_back('\b'); break;
          case 'f': s->push_back('\f'); break;
          case 'u': {
            // keep ASCII, replace the rest by '?': config keys/values on the path are ASCII
            unsigned code = 0;
            for (int i = 0; i < 4; ++i) {
              ++p_;
              if (!*p_) return fail("bad \\u escape");
              code = code * 16 + static_cast<unsigned>(strtol(std::string(1, *p_).c_str(), nullptr, 16));
            }
            s->push_back(code < 128 ? static_cast<char>(code) : '?');
            break;
          }
          case '\0': return fail("unterminated string");
          default: s->push_back(*p_);
        }
        ++p_;
      } else {
        s->push_back(*p_++);
      }
    }
    if (*p_ != '"') return fail("unterminated string");
    ++p_;
    return true;
  }
  bool value(JValue* v) {
    ws();
    if (*p_ == '{') {
      v->type = JValue::Obj;
      ++p_;
      ws();
      if (*p_ == '}') { ++p_; return true; }
      while (true) {
        ws();
        std::string k;
        if (!string(&k)) return false;
        ws();
        if (*p_ != ':') return fail("expected ':'");
        ++p_;
        JValue child;
        if (!value(&child)) return false;
        v->obj.emplace_back(std::move(k), std::move(child));
        ws();
        if (*p_ == ',') { ++p_; continue; }
        if (*p_ == '}') { ++p_; return true; }
        return fail("expected ',' or '}'");
      }
    }
    if (*p_ == '[') {
      v->type = JValue::Arr;
      ++p_;
      ws();
      if (*p_ == ']') { ++p_; return true; }
      while (true) {
        JValue child;
        if (!value(&child)) return false;
        v->arr.push_back(std::move(child));
        ws();
        if (*p_ == ',') { ++p_; continue; }
        if (*p_ == ']') { ++p_; return true; }
        return fail("expected ',' or ']'");
      }
    }
    if (*p_ == '"') {
      v->type = JValue::Str;
      return string(&v->str);
    }
    if (!strncmp(p_, "true", 4)) { v->type = JValue::Bool; v->b = true; p_ += 4; return true; }
    if (!strncmp(p_, "false", 5)) { v->type = JValue::Bool; v->b = false; p_ += 5; return true; }
    if (!strncmp(p_, "null", 4)) { v->type = JValue::Null; p_ += 4; return true; }
    char* end = nullptr;
    const double d = strtod(p_, &end);
    if (end == p_) return fail("unexpected character");
    v->type = JValue::Num;
    v->num = d;
    p_ = end;
    return true;
  }
};

}  // namespace xv
