// Minimal JSON reader for the experiment configs (workspace/experiments/config/*.json in the
// reference). Objects, arrays, numbers, strings, true/false/null — enough for base_config.json
// and the task files; no dependency (the reference uses nlohmann/json, absent here).
#pragma once
#include <cctype>
#include <cstdlib>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace mpccbf {

struct JsonValue {
    enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<JsonValue> arr;
    std::map<std::string, JsonValue> obj;

    const JsonValue& at(const std::string& key) const {
        if (kind != Object) throw std::runtime_error("json: not an object while looking up '" + key + "'");
        auto it = obj.find(key);
        if (it == obj.end()) throw std::runtime_error("json: missing key '" + key + "'");
        return it->second;
    }
    bool has(const std::string& key) const { return kind == Object && obj.count(key) > 0; }
    const JsonValue& at(size_t i) const {
        if (kind != Array || i >= arr.size()) throw std::runtime_error("json: array index out of range");
        return arr[i];
    }
    double number() const {
        if (kind == Number) return num;
        if (kind == Bool) return b ? 1.0 : 0.0;
        throw std::runtime_error("json: expected a number");
    }
    bool boolean() const {
        if (kind == Bool) return b;
        if (kind == Number) return num != 0.0;
        throw std::runtime_error("json: expected a boolean");
    }
    size_t size() const { return kind == Array ? arr.size() : obj.size(); }
};

class JsonParser {
public:
    explicit JsonParser(const std::string& text) : s_(text) {}
    JsonValue parse() {
        JsonValue v = value();
        ws();
        if (i_ != s_.size()) fail("trailing characters");
        return v;
    }

private:
    const std::string& s_;
    size_t i_ = 0;
    [[noreturn]] void fail(const std::string& m) const {
        throw std::runtime_error("json parse error at byte " + std::to_string(i_) + ": " + m);
    }
    void ws() {
        while (i_ < s_.size() && std::isspace(static_cast<unsigned char>(s_[i_]))) ++i_;
    }
    bool eat(char c) {
        ws();
        if (i_ < s_.size() && s_[i_] == c) {
            ++i_;
            return true;
        }
        return false;
    }
    JsonValue value() {
        ws();
        if (i_ >= s_.size()) fail("unexpected end");
        const char c = s_[i_];
        JsonValue v;
        if (c == '{') {
            ++i_;
            v.kind = JsonValue::Object;
            if (eat('}')) return v;
            do {
                ws();
                const std::string key = string();
                if (!eat(':')) fail("expected ':'");
                v.obj[key] = value();
            } while (eat(','));
            if (!eat('}')) fail("expected '}'");
        } else if (c == '[') {
            ++i_;
            v.kind = JsonValue::Array;
            if (eat(']')) return v;
            do {
                v.arr.push_back(value());
            } while (eat(','));
            if (!eat(']')) fail("expected ']'");
        } else if (c == '"') {
            v.kind = JsonValue::String;
            v.str = string();
        } else if (s_.compare(i_, 4, "true") == 0) {
            v.kind = JsonValue::Bool;
            v.b = true;
            i_ += 4;
        } else if (s_.compare(i_, 5, "false") == 0) {
            v.kind = JsonValue::Bool;
            v.b = false;
            i_ += 5;
        } else if (s_.compare(i_, 4, "null") == 0) {
            i_ += 4;
        } else {
            const char* begin = s_.c_str() + i_;
            char* end = nullptr;
            v.num = std::strtod(begin, &end);
            if (end == begin) fail("unexpected character");
            v.kind = JsonValue::Number;
            i_ += static_cast<size_t>(end - begin);
        }
        return v;
    }
    std::string string() {
        if (i_ >= s_.size() || s_[i_] != '"') fail("expected string");
        ++i_;
        std::string out;
        while (i_ < s_.size() && s_[i_] != '"') {
            if (s_[i_] == '\\' && i_ + 1 < s_.size()) {
                ++i_;
                switch (s_[i_]) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': i_ += 4; out += '?'; break;
                    default: out += s_[i_];
                }
                ++i_;
            } else {
                out += s_[i_++];
            }
        }
        if (i_ >= s_.size()) fail("unterminated string");
        ++i_;
        return out;
    }
};

}  // namespace mpccbf
