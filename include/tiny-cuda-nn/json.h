/* tcnn::json - the JSON value type of the facade.
 *
 * The reference's API speaks nlohmann::json (`using json = nlohmann::json;`, include/tiny-cuda-nn/common.h:38-43 via
 * dependencies/json/json.hpp): create_from_config(n_in, n_out, json), Trainer::serialize() -> json with binary blobs
 * (trainer.h:275-315, gpu_memory_json.h:36-72), hyperparams() -> json, update_hyperparams(json).  This header provides the
 * subset of that interface the hot path's hosts use, written from scratch (no third-party code): brace-initialiser
 * construction with nlohmann's object/array rule, value(key, default), contains, operator[], at, get<T> and implicit
 * conversion, items of arrays/objects, parse (with // and block comments when asked, as the samples do), dump, binary values, and
 * to_msgpack / from_msgpack so that snapshots travel between hosts of the reference and hosts of this library.
 * A host that already uses nlohmann::json can keep it: every entry point that takes a tcnn::json also takes any type with
 * a dump() method. */
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <initializer_list>
#include <istream>
#include <iterator>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

namespace tcnn {

class json {
public:
	enum class value_t : uint8_t { null, object, array, string, boolean, number_integer, number_unsigned, number_float, binary };
	using object_t = std::map<std::string, json>;
	using array_t = std::vector<json>;
	struct binary_t : std::vector<uint8_t> { /* nlohmann::byte_container_with_subtype: bytes + optional subtype */
		using std::vector<uint8_t>::vector;
		binary_t() = default;
		binary_t(const std::vector<uint8_t>& v) : std::vector<uint8_t>(v) {}
		binary_t(std::vector<uint8_t>&& v) : std::vector<uint8_t>(std::move(v)) {}
		bool has_subtype() const { return m_has_subtype; }
		uint64_t subtype() const { return m_subtype; }
		void set_subtype(uint64_t s) { m_subtype = s; m_has_subtype = true; }
	private:
		uint64_t m_subtype = 0; bool m_has_subtype = false;
	};

	/* ---- construction ---- */
	json() = default;
	json(std::nullptr_t) {}
	json(bool v) : m_type(value_t::boolean) { m_num.b = v; }
	template <typename T, typename std::enable_if<std::is_integral<T>::value && std::is_signed<T>::value && !std::is_same<T, bool>::value, int>::type = 0>
	json(T v) : m_type(value_t::number_integer) { m_num.i = (int64_t)v; }
	template <typename T, typename std::enable_if<std::is_integral<T>::value && std::is_unsigned<T>::value && !std::is_same<T, bool>::value, int>::type = 0>
	json(T v) : m_type(value_t::number_unsigned) { m_num.u = (uint64_t)v; }
	template <typename T, typename std::enable_if<std::is_floating_point<T>::value, int>::type = 0>
	json(T v) : m_type(value_t::number_float) { m_num.f = (double)v; }
	json(const char* s) : m_type(value_t::string), m_str(std::make_shared<std::string>(s)) {}
	json(const std::string& s) : m_type(value_t::string), m_str(std::make_shared<std::string>(s)) {}
	json(const binary_t& b) : m_type(value_t::binary), m_bin(std::make_shared<binary_t>(b)) {}
	json(binary_t&& b) : m_type(value_t::binary), m_bin(std::make_shared<binary_t>(std::move(b))) {}
	json(const object_t& o) : m_type(value_t::object), m_obj(std::make_shared<object_t>(o)) {}
	json(const array_t& a) : m_type(value_t::array), m_arr(std::make_shared<array_t>(a)) {}
	template <typename T, typename std::enable_if<std::is_arithmetic<T>::value, int>::type = 0>
	json(const std::vector<T>& v) : m_type(value_t::array), m_arr(std::make_shared<array_t>(v.begin(), v.end())) {}
	json(const std::vector<std::string>& v) : m_type(value_t::array), m_arr(std::make_shared<array_t>(v.begin(), v.end())) {}
	/* nlohmann's rule: a brace list whose elements are all [string, value] pairs is an object, anything else an array */
	json(std::initializer_list<json> init) {
		bool is_object = true;
		for (const json& e : init) is_object = is_object && e.is_array() && e.size() == 2 && (*e.m_arr)[0].is_string();
		if (is_object) {
			m_type = value_t::object; m_obj = std::make_shared<object_t>();
			for (const json& e : init) (*m_obj)[*(*e.m_arr)[0].m_str] = (*e.m_arr)[1];
		} else {
			m_type = value_t::array; m_arr = std::make_shared<array_t>(init.begin(), init.end());
		}
	}
	json(const json& o) { copy_from(o); }
	json(json&& o) noexcept = default;
	json& operator=(const json& o) { if (this != &o) { json t(o); *this = std::move(t); } return *this; }
	json& operator=(json&& o) noexcept = default;

	static json object() { json j; j.m_type = value_t::object; j.m_obj = std::make_shared<object_t>(); return j; }
	static json array() { json j; j.m_type = value_t::array; j.m_arr = std::make_shared<array_t>(); return j; }
	static json binary(const std::vector<uint8_t>& bytes) { return json(binary_t(bytes)); }

	/* ---- inspection ---- */
	value_t type() const { return m_type; }
	bool is_null() const { return m_type == value_t::null; }
	bool is_object() const { return m_type == value_t::object; }
	bool is_array() const { return m_type == value_t::array; }
	bool is_string() const { return m_type == value_t::string; }
	bool is_boolean() const { return m_type == value_t::boolean; }
	bool is_number() const { return m_type == value_t::number_integer || m_type == value_t::number_unsigned || m_type == value_t::number_float; }
	bool is_number_integer() const { return m_type == value_t::number_integer || m_type == value_t::number_unsigned; }
	bool is_number_float() const { return m_type == value_t::number_float; }
	bool is_binary() const { return m_type == value_t::binary; }
	bool empty() const { return is_null() || (is_object() && m_obj->empty()) || (is_array() && m_arr->empty()); }
	size_t size() const { return is_object() ? m_obj->size() : is_array() ? m_arr->size() : is_null() ? 0 : 1; }
	bool contains(const std::string& key) const { return is_object() && m_obj->count(key) != 0; }

	/* ---- access ---- */
	json& operator[](const std::string& key) {
		if (is_null()) { m_type = value_t::object; m_obj = std::make_shared<object_t>(); }
		if (!is_object()) throw std::runtime_error("json: operator[] with a string key on a value that is not an object");
		return (*m_obj)[key];
	}
	json& operator[](const char* key) { return (*this)[std::string(key)]; }
	const json& operator[](const std::string& key) const { return at(key); }
	const json& operator[](const char* key) const { return at(std::string(key)); }
	json& operator[](size_t i) { if (!is_array() || i >= m_arr->size()) throw std::runtime_error("json: array index out of range"); return (*m_arr)[i]; }
	const json& operator[](size_t i) const { if (!is_array() || i >= m_arr->size()) throw std::runtime_error("json: array index out of range"); return (*m_arr)[i]; }
	json& operator[](int i) { return (*this)[(size_t)i]; }
	const json& operator[](int i) const { return (*this)[(size_t)i]; }
	const json& at(const std::string& key) const {
		if (!is_object()) throw std::runtime_error("json: at('" + key + "') on a value that is not an object");
		auto it = m_obj->find(key);
		if (it == m_obj->end()) throw std::runtime_error("json: key '" + key + "' not found");
		return it->second;
	}
	json& at(const std::string& key) { return const_cast<json&>(static_cast<const json&>(*this).at(key)); }
	void push_back(const json& v) {
		if (is_null()) { m_type = value_t::array; m_arr = std::make_shared<array_t>(); }
		if (!is_array()) throw std::runtime_error("json: push_back on a value that is not an array");
		m_arr->push_back(v);
	}
	void erase(const std::string& key) { if (is_object()) m_obj->erase(key); }
	const object_t& items() const { if (!is_object()) throw std::runtime_error("json: items() on a value that is not an object"); return *m_obj; }
	const array_t& elements() const { if (!is_array()) throw std::runtime_error("json: elements() on a value that is not an array"); return *m_arr; }
	array_t::const_iterator begin() const { return elements().begin(); }
	array_t::const_iterator end() const { return elements().end(); }
	const binary_t& get_binary() const { if (!is_binary()) throw std::runtime_error("json: get_binary() on a value that is not binary"); return *m_bin; }
	binary_t& get_binary() { if (!is_binary()) throw std::runtime_error("json: get_binary() on a value that is not binary"); return *m_bin; }

	/* value(key, default): the default's type decides the result type (a string literal default gives std::string) */
	template <typename T>
	T value(const std::string& key, const T& dflt) const { return contains(key) ? at(key).template get<T>() : dflt; }
	std::string value(const std::string& key, const char* dflt) const { return contains(key) ? at(key).get<std::string>() : std::string(dflt); }

	template <typename T> T get() const { return get_impl(static_cast<T*>(nullptr)); }
	template <typename T> void get_to(T& out) const { out = get<T>(); }
	/* implicit conversion like nlohmann's: `float lr = j["learning_rate"];`, `std::string s = j["otype"];` */
	template <typename T, typename std::enable_if<std::is_arithmetic<T>::value || std::is_same<T, std::string>::value, int>::type = 0>
	operator T() const { return get<T>(); }

	bool operator==(const json& o) const {
		if (is_number() && o.is_number()) return as_double() == o.as_double();
		if (m_type != o.m_type) return false;
		switch (m_type) {
			case value_t::null: return true;
			case value_t::boolean: return m_num.b == o.m_num.b;
			case value_t::string: return *m_str == *o.m_str;
			case value_t::binary: return static_cast<const std::vector<uint8_t>&>(*m_bin) == static_cast<const std::vector<uint8_t>&>(*o.m_bin);
			case value_t::array: return *m_arr == *o.m_arr;
			case value_t::object: return *m_obj == *o.m_obj;
			default: return false;
		}
	}
	bool operator!=(const json& o) const { return !(*this == o); }

	/* ---- text ---- */
	std::string dump(int indent = -1) const { std::string out; dump_to(out, indent, 0); return out; }
	static json parse(const std::string& text, std::nullptr_t = nullptr, bool allow_exceptions = true, bool ignore_comments = false) {
		(void)allow_exceptions;
		Reader r{text, 0, ignore_comments};
		json v = r.value();
		r.skip();
		if (r.pos != text.size()) r.fail("trailing characters");
		return v;
	}
	static json parse(std::istream& in, std::nullptr_t = nullptr, bool allow_exceptions = true, bool ignore_comments = false) {
		std::string text((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
		return parse(text, nullptr, allow_exceptions, ignore_comments);
	}
	friend std::ostream& operator<<(std::ostream& os, const json& j) { return os << j.dump(); }

	/* ---- MessagePack (the wire form of snapshots: json::to_msgpack(trainer->serialize())) ---- */
	static std::vector<uint8_t> to_msgpack(const json& j) { std::vector<uint8_t> out; j.pack(out); return out; }
	static json from_msgpack(const std::vector<uint8_t>& bytes) {
		size_t pos = 0;
		json v = unpack(bytes.data(), bytes.size(), pos);
		if (pos != bytes.size()) throw std::runtime_error("json::from_msgpack: trailing bytes");
		return v;
	}
	static json from_msgpack(const uint8_t* data, size_t n) { size_t pos = 0; return unpack(data, n, pos); }

private:
	value_t m_type = value_t::null;
	union { bool b; int64_t i; uint64_t u; double f; } m_num = {};
	std::shared_ptr<std::string> m_str;
	std::shared_ptr<binary_t> m_bin;
	std::shared_ptr<array_t> m_arr;
	std::shared_ptr<object_t> m_obj;

	void copy_from(const json& o) { /* deep copy: values, not references */
		m_type = o.m_type; m_num = o.m_num;
		if (o.m_str) m_str = std::make_shared<std::string>(*o.m_str);
		if (o.m_bin) m_bin = std::make_shared<binary_t>(*o.m_bin);
		if (o.m_arr) m_arr = std::make_shared<array_t>(*o.m_arr);
		if (o.m_obj) m_obj = std::make_shared<object_t>(*o.m_obj);
	}
	double as_double() const {
		switch (m_type) {
			case value_t::number_integer: return (double)m_num.i;
			case value_t::number_unsigned: return (double)m_num.u;
			case value_t::number_float: return m_num.f;
			case value_t::boolean: return m_num.b ? 1.0 : 0.0;
			default: throw std::runtime_error("json: value is not a number (" + dump() + ")");
		}
	}
	template <typename T, typename std::enable_if<std::is_arithmetic<T>::value && !std::is_same<T, bool>::value, int>::type = 0>
	T get_impl(T*) const {
		if (m_type == value_t::number_integer) return (T)m_num.i;
		if (m_type == value_t::number_unsigned) return (T)m_num.u;
		return (T)as_double();
	}
	bool get_impl(bool*) const {
		if (m_type == value_t::boolean) return m_num.b;
		if (is_number()) return as_double() != 0.0;
		throw std::runtime_error("json: value is not a boolean (" + dump() + ")");
	}
	std::string get_impl(std::string*) const { if (!is_string()) throw std::runtime_error("json: value is not a string (" + dump() + ")"); return *m_str; }
	json get_impl(json*) const { return *this; }
	binary_t get_impl(binary_t*) const { return get_binary(); }
	template <typename T> std::vector<T> get_impl(std::vector<T>*) const {
		std::vector<T> v;
		for (const json& e : elements()) v.push_back(e.template get<T>());
		return v;
	}

	/* ---- writer ---- */
	static void escape_to(std::string& out, const std::string& s) {
		out += '"';
		for (unsigned char c : s) {
			switch (c) {
				case '"': out += "\\\""; break; case '\\': out += "\\\\"; break; case '\n': out += "\\n"; break; case '\t': out += "\\t"; break;
				case '\r': out += "\\r"; break; case '\b': out += "\\b"; break; case '\f': out += "\\f"; break;
				default: if (c < 0x20) { char b[8]; std::snprintf(b, sizeof(b), "\\u%04x", c); out += b; } else out += (char)c;
			}
		}
		out += '"';
	}
	void dump_to(std::string& out, int indent, int depth) const {
		const bool pretty = indent >= 0;
		auto newline = [&](int d) { if (pretty) { out += '\n'; out.append((size_t)(d * indent), ' '); } };
		switch (m_type) {
			case value_t::null: out += "null"; break;
			case value_t::boolean: out += m_num.b ? "true" : "false"; break;
			case value_t::number_integer: out += std::to_string(m_num.i); break;
			case value_t::number_unsigned: out += std::to_string(m_num.u); break;
			case value_t::number_float: {
				if (!std::isfinite(m_num.f)) { out += "null"; break; }
				char b[40]; std::snprintf(b, sizeof(b), "%.17g", m_num.f);
				/* shortest representation that round-trips, and always recognisable as a float */
				for (int p = 1; p < 17; ++p) { char t[40]; std::snprintf(t, sizeof(t), "%.*g", p, m_num.f); if (std::strtod(t, nullptr) == m_num.f) { std::strcpy(b, t); break; } }
				out += b;
				if (!std::strpbrk(b, ".eE")) out += ".0";
				break;
			}
			case value_t::string: escape_to(out, *m_str); break;
			case value_t::binary: { /* nlohmann's text form of a binary value: {"bytes": [...], "subtype": null} */
				out += pretty ? "{" : "{"; newline(depth + 1); out += "\"bytes\":"; if (pretty) out += ' '; out += '[';
				for (size_t i = 0; i < m_bin->size(); ++i) { if (i) out += ','; out += std::to_string((unsigned)(*m_bin)[i]); }
				out += "],"; newline(depth + 1); out += "\"subtype\":"; if (pretty) out += ' ';
				out += m_bin->has_subtype() ? std::to_string(m_bin->subtype()) : "null"; newline(depth); out += '}';
				break;
			}
			case value_t::array: {
				if (m_arr->empty()) { out += "[]"; break; }
				out += '[';
				for (size_t i = 0; i < m_arr->size(); ++i) { if (i) out += ','; newline(depth + 1); (*m_arr)[i].dump_to(out, indent, depth + 1); }
				newline(depth); out += ']';
				break;
			}
			case value_t::object: {
				if (m_obj->empty()) { out += "{}"; break; }
				out += '{';
				bool first = true;
				for (const auto& kv : *m_obj) {
					if (!first) out += ',';
					first = false;
					newline(depth + 1); escape_to(out, kv.first); out += ':'; if (pretty) out += ' ';
					kv.second.dump_to(out, indent, depth + 1);
				}
				newline(depth); out += '}';
				break;
			}
		}
	}

	/* ---- reader ---- */
	struct Reader {
		const std::string& s; size_t pos; bool comments;
		[[noreturn]] void fail(const std::string& m) const { throw std::runtime_error("json: parse error at offset " + std::to_string(pos) + ": " + m); }
		void skip() {
			for (;;) {
				while (pos < s.size() && (s[pos] == ' ' || s[pos] == '\n' || s[pos] == '\t' || s[pos] == '\r')) ++pos;
				if (comments && pos + 1 < s.size() && s[pos] == '/' && s[pos + 1] == '/') { while (pos < s.size() && s[pos] != '\n') ++pos; continue; }
				if (comments && pos + 1 < s.size() && s[pos] == '/' && s[pos + 1] == '*') {
					const size_t e = s.find("*/", pos + 2);
					if (e == std::string::npos) fail("unterminated comment");
					pos = e + 2; continue;
				}
				return;
			}
		}
		char peek() { skip(); if (pos >= s.size()) fail("unexpected end of input"); return s[pos]; }
		bool literal(const char* w) { const size_t n = std::strlen(w); if (s.compare(pos, n, w) == 0) { pos += n; return true; } return false; }
		std::string string() {
			if (peek() != '"') fail("expected a string");
			++pos;
			std::string r;
			while (pos < s.size() && s[pos] != '"') {
				char c = s[pos++];
				if (c != '\\') { r += c; continue; }
				if (pos >= s.size()) fail("bad escape");
				const char e = s[pos++];
				switch (e) {
					case 'n': r += '\n'; break; case 't': r += '\t'; break; case 'r': r += '\r'; break; case 'b': r += '\b'; break; case 'f': r += '\f'; break;
					case 'u': {
						if (pos + 4 > s.size()) fail("bad \\u escape");
						const unsigned cp = (unsigned)std::strtoul(s.substr(pos, 4).c_str(), nullptr, 16); pos += 4;
						if (cp < 0x80) r += (char)cp;
						else if (cp < 0x800) { r += (char)(0xC0 | (cp >> 6)); r += (char)(0x80 | (cp & 0x3F)); }
						else { r += (char)(0xE0 | (cp >> 12)); r += (char)(0x80 | ((cp >> 6) & 0x3F)); r += (char)(0x80 | (cp & 0x3F)); }
						break;
					}
					default: r += e;
				}
			}
			if (pos >= s.size()) fail("unterminated string");
			++pos;
			return r;
		}
		json value() {
			const char c = peek();
			if (c == '{') {
				++pos;
				json o = json::object();
				if (peek() == '}') { ++pos; return o; }
				for (;;) {
					std::string k = string();
					if (peek() != ':') fail("expected ':'");
					++pos;
					(*o.m_obj)[k] = value();
					const char d = peek(); ++pos;
					if (d == '}') break;
					if (d != ',') fail("expected ',' or '}'");
				}
				return o;
			}
			if (c == '[') {
				++pos;
				json a = json::array();
				if (peek() == ']') { ++pos; return a; }
				for (;;) {
					a.m_arr->push_back(value());
					const char d = peek(); ++pos;
					if (d == ']') break;
					if (d != ',') fail("expected ',' or ']'");
				}
				return a;
			}
			if (c == '"') return json(string());
			if (literal("true")) return json(true);
			if (literal("false")) return json(false);
			if (literal("null")) return json();
			const char* b = s.c_str() + pos; char* e = nullptr;
			const double d = std::strtod(b, &e);
			if (e == b) fail("unexpected character");
			const std::string tok(b, (size_t)(e - b));
			pos += tok.size();
			if (tok.find_first_of(".eE") == std::string::npos) { /* integer literal: keep it exact */
				if (tok[0] == '-') return json((int64_t)std::strtoll(tok.c_str(), nullptr, 10));
				return json((uint64_t)std::strtoull(tok.c_str(), nullptr, 10));
			}
			return json(d);
		}
	};

	/* ---- MessagePack ---- */
	static void be(std::vector<uint8_t>& out, uint64_t v, int bytes) { for (int i = bytes - 1; i >= 0; --i) out.push_back((uint8_t)(v >> (8 * i))); }
	void pack(std::vector<uint8_t>& out) const {
		switch (m_type) {
			case value_t::null: out.push_back(0xC0); break;
			case value_t::boolean: out.push_back(m_num.b ? 0xC3 : 0xC2); break;
			case value_t::number_integer:
				if (m_num.i >= 0) { pack_unsigned(out, (uint64_t)m_num.i); break; }
				if (m_num.i >= -32) out.push_back((uint8_t)m_num.i);
				else if (m_num.i >= INT8_MIN) { out.push_back(0xD0); be(out, (uint64_t)m_num.i, 1); }
				else if (m_num.i >= INT16_MIN) { out.push_back(0xD1); be(out, (uint64_t)m_num.i, 2); }
				else if (m_num.i >= INT32_MIN) { out.push_back(0xD2); be(out, (uint64_t)m_num.i, 4); }
				else { out.push_back(0xD3); be(out, (uint64_t)m_num.i, 8); }
				break;
			case value_t::number_unsigned: pack_unsigned(out, m_num.u); break;
			case value_t::number_float: {
				const float f = (float)m_num.f;
				if ((double)f == m_num.f) { uint32_t u; std::memcpy(&u, &f, 4); out.push_back(0xCA); be(out, u, 4); } /* nlohmann writes float32 when lossless */
				else { uint64_t u; std::memcpy(&u, &m_num.f, 8); out.push_back(0xCB); be(out, u, 8); }
				break;
			}
			case value_t::string: {
				const size_t n = m_str->size();
				if (n <= 31) out.push_back((uint8_t)(0xA0 | n));
				else if (n <= 0xFF) { out.push_back(0xD9); be(out, n, 1); }
				else if (n <= 0xFFFF) { out.push_back(0xDA); be(out, n, 2); }
				else { out.push_back(0xDB); be(out, n, 4); }
				out.insert(out.end(), m_str->begin(), m_str->end());
				break;
			}
			case value_t::binary: {
				const size_t n = m_bin->size();
				if (n <= 0xFF) { out.push_back(0xC4); be(out, n, 1); }
				else if (n <= 0xFFFF) { out.push_back(0xC5); be(out, n, 2); }
				else if (n <= 0xFFFFFFFFull) { out.push_back(0xC6); be(out, n, 4); }
				else throw std::runtime_error("json::to_msgpack: binary value larger than 4 GiB");
				out.insert(out.end(), m_bin->begin(), m_bin->end());
				break;
			}
			case value_t::array: {
				const size_t n = m_arr->size();
				if (n <= 15) out.push_back((uint8_t)(0x90 | n));
				else if (n <= 0xFFFF) { out.push_back(0xDC); be(out, n, 2); }
				else { out.push_back(0xDD); be(out, n, 4); }
				for (const json& e : *m_arr) e.pack(out);
				break;
			}
			case value_t::object: {
				const size_t n = m_obj->size();
				if (n <= 15) out.push_back((uint8_t)(0x80 | n));
				else if (n <= 0xFFFF) { out.push_back(0xDE); be(out, n, 2); }
				else { out.push_back(0xDF); be(out, n, 4); }
				for (const auto& kv : *m_obj) { json(kv.first).pack(out); kv.second.pack(out); }
				break;
			}
		}
	}
	static void pack_unsigned(std::vector<uint8_t>& out, uint64_t u) {
		if (u <= 0x7F) out.push_back((uint8_t)u);
		else if (u <= 0xFF) { out.push_back(0xCC); be(out, u, 1); }
		else if (u <= 0xFFFF) { out.push_back(0xCD); be(out, u, 2); }
		else if (u <= 0xFFFFFFFFull) { out.push_back(0xCE); be(out, u, 4); }
		else { out.push_back(0xCF); be(out, u, 8); }
	}
	static uint64_t rd(const uint8_t* d, size_t n, size_t& pos, int bytes) {
		if (pos + (size_t)bytes > n) throw std::runtime_error("json::from_msgpack: unexpected end of input");
		uint64_t v = 0;
		for (int i = 0; i < bytes; ++i) v = (v << 8) | d[pos++];
		return v;
	}
	static json unpack(const uint8_t* d, size_t n, size_t& pos) {
		const uint8_t t = (uint8_t)rd(d, n, pos, 1);
		auto bytes = [&](size_t len) { if (pos + len > n) throw std::runtime_error("json::from_msgpack: unexpected end of input"); const uint8_t* p = d + pos; pos += len; return p; };
		auto arr = [&](size_t len) { json a = json::array(); for (size_t i = 0; i < len; ++i) a.m_arr->push_back(unpack(d, n, pos)); return a; };
		auto map = [&](size_t len) {
			json o = json::object();
			for (size_t i = 0; i < len; ++i) { json k = unpack(d, n, pos); if (!k.is_string()) throw std::runtime_error("json::from_msgpack: map key is not a string"); (*o.m_obj)[*k.m_str] = unpack(d, n, pos); }
			return o;
		};
		auto str = [&](size_t len) { const uint8_t* p = bytes(len); return json(std::string((const char*)p, len)); };
		auto bin = [&](size_t len) { const uint8_t* p = bytes(len); return json(binary_t(std::vector<uint8_t>(p, p + len))); };
		if (t <= 0x7F) return json((uint64_t)t);
		if (t >= 0xE0) return json((int64_t)(int8_t)t);
		if ((t & 0xF0) == 0x80) return map(t & 0x0F);
		if ((t & 0xF0) == 0x90) return arr(t & 0x0F);
		if ((t & 0xE0) == 0xA0) return str(t & 0x1F);
		switch (t) {
			case 0xC0: return json();
			case 0xC2: return json(false);
			case 0xC3: return json(true);
			case 0xC4: return bin((size_t)rd(d, n, pos, 1));
			case 0xC5: return bin((size_t)rd(d, n, pos, 2));
			case 0xC6: return bin((size_t)rd(d, n, pos, 4));
			case 0xCA: { const uint32_t u = (uint32_t)rd(d, n, pos, 4); float f; std::memcpy(&f, &u, 4); return json((double)f); }
			case 0xCB: { const uint64_t u = rd(d, n, pos, 8); double f; std::memcpy(&f, &u, 8); return json(f); }
			case 0xCC: return json((uint64_t)rd(d, n, pos, 1));
			case 0xCD: return json((uint64_t)rd(d, n, pos, 2));
			case 0xCE: return json((uint64_t)rd(d, n, pos, 4));
			case 0xCF: return json((uint64_t)rd(d, n, pos, 8));
			case 0xD0: return json((int64_t)(int8_t)rd(d, n, pos, 1));
			case 0xD1: return json((int64_t)(int16_t)rd(d, n, pos, 2));
			case 0xD2: return json((int64_t)(int32_t)rd(d, n, pos, 4));
			case 0xD3: return json((int64_t)rd(d, n, pos, 8));
			case 0xD9: return str((size_t)rd(d, n, pos, 1));
			case 0xDA: return str((size_t)rd(d, n, pos, 2));
			case 0xDB: return str((size_t)rd(d, n, pos, 4));
			case 0xDC: return arr((size_t)rd(d, n, pos, 2));
			case 0xDD: return arr((size_t)rd(d, n, pos, 4));
			case 0xDE: return map((size_t)rd(d, n, pos, 2));
			case 0xDF: return map((size_t)rd(d, n, pos, 4));
			default: throw std::runtime_error("json::from_msgpack: unsupported type byte " + std::to_string((unsigned)t));
		}
	}
};

} // namespace tcnn
