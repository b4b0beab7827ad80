// Minimal JSON reader for the tcnn config objects (objects, arrays, strings, numbers, true/false/null).
// Only what plan_create needs: nested objects with case-sensitive keys and case-insensitive otype strings
// (the reference dispatches otype through a case-insensitive map, include/tiny-cuda-nn/common_host.h:242-246).
#pragma once
#include <cctype>
#include <cmath>
#include <cstdlib>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace jsonmini {

struct Value {
	enum Kind { Null, Bool, Number, String, Array, Object } kind = Null;
	bool b = false;
	double num = 0;
	std::string str;
	std::vector<Value> arr;
	std::map<std::string, Value> obj;

	bool contains(const std::string& k) const { return kind == Object && obj.count(k) != 0; }
	std::vector<std::string> keys() const { std::vector<std::string> r; for (const auto& kv : obj) r.push_back(kv.first); return r; }
	const Value& at(const std::string& k) const {
		auto it = obj.find(k);
		if (it == obj.end()) throw std::runtime_error("json: missing key '" + k + "'");
		return it->second;
	}
	double number(const std::string& k, double dflt) const {
		if (!contains(k)) return dflt;
		const Value& v = at(k);
		if (v.kind == Number) return v.num;
		if (v.kind == Bool) return v.b ? 1.0 : 0.0;
		throw std::runtime_error("json: key '" + k + "' is not a number");
	}
	bool boolean(const std::string& k, bool dflt) const {
		if (!contains(k)) return dflt;
		const Value& v = at(k);
		if (v.kind == Bool) return v.b;
		if (v.kind == Number) return v.num != 0;
		throw std::runtime_error("json: key '" + k + "' is not a bool");
	}
	std::string string(const std::string& k, const std::string& dflt) const {
		if (!contains(k)) return dflt;
		const Value& v = at(k);
		if (v.kind != String) throw std::runtime_error("json: key '" + k + "' is not a string");
		return v.str;
	}
	Value object(const std::string& k) const {
		if (!contains(k)) { Value v; v.kind = Object; return v; }
		return at(k);
	}
};

class Parser {
public:
	explicit Parser(const std::string& s) : s_(s) {}
	Value parse() {
		Value v = value();
		ws();
		if (i_ != s_.size()) fail("trailing characters");
		return v;
	}

private:
	const std::string& s_;
	size_t i_ = 0;
	[[noreturn]] void fail(const std::string& m) { throw std::runtime_error("json: " + m + " at offset " + std::to_string(i_)); }
	void ws() { while (i_ < s_.size() && std::isspace((unsigned char)s_[i_])) ++i_; }
	char peek() { ws(); if (i_ >= s_.size()) fail("unexpected end"); return s_[i_]; }
	void expect(char c) { if (peek() != c) fail(std::string("expected '") + c + "'"); ++i_; }
	bool lit(const char* w) { size_t n = std::char_traits<char>::length(w); if (s_.compare(i_, n, w) == 0) { i_ += n; return true; } return false; }
	Value value() {
		char c = peek();
		Value v;
		if (c == '{') {
			v.kind = Value::Object; ++i_;
			if (peek() == '}') { ++i_; return v; }
			for (;;) {
				if (peek() != '"') fail("expected string key");
				std::string k = str();
				expect(':');
				v.obj[k] = value();
				char d = peek(); ++i_;
				if (d == '}') break;
				if (d != ',') fail("expected ',' or '}'");
			}
		} else if (c == '[') {
			v.kind = Value::Array; ++i_;
			if (peek() == ']') { ++i_; return v; }
			for (;;) {
				v.arr.push_back(value());
				char d = peek(); ++i_;
				if (d == ']') break;
				if (d != ',') fail("expected ',' or ']'");
			}
		} else if (c == '"') {
			v.kind = Value::String; v.str = str();
		} else if (lit("true")) { v.kind = Value::Bool; v.b = true;
		} else if (lit("false")) { v.kind = Value::Bool; v.b = false;
		} else if (lit("null")) { v.kind = Value::Null;
		} else {
			const char* b = s_.c_str() + i_; char* e = nullptr;
			v.num = std::strtod(b, &e);
			if (e == b) fail("bad value");
			v.kind = Value::Number; i_ += (size_t)(e - b);
		}
		return v;
	}
	std::string str() {
		expect('"');
		std::string r;
		while (i_ < s_.size() && s_[i_] != '"') {
			char c = s_[i_++];
			if (c == '\\') {
				if (i_ >= s_.size()) fail("bad escape");
				char e = s_[i_++];
				switch (e) {
					case 'n': r += '\n'; break; case 't': r += '\t'; break; case 'r': r += '\r'; break;
					case 'b': r += '\b'; break; case 'f': r += '\f'; break;
					case 'u': { if (i_ + 4 > s_.size()) fail("bad \\u"); unsigned cp = (unsigned)std::strtoul(s_.substr(i_, 4).c_str(), nullptr, 16); i_ += 4; r += (cp < 128) ? (char)cp : '?'; break; }
					default: r += e;
				}
			} else r += c;
		}
		if (i_ >= s_.size()) fail("unterminated string");
		++i_;
		return r;
	}
};

inline Value parse(const std::string& s) { return Parser(s).parse(); }
inline std::string lower(std::string s) { for (auto& c : s) c = (char)std::tolower((unsigned char)c); return s; }

} // namespace jsonmini
