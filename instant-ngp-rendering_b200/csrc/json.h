// Minimal JSON value + recursive-descent reader/writer for the tcnn / instant-ngp config
// schema. Host only. Accepts `//` and `/* */` comments, which the reference's config
// files use (configs/image/base.json:6-9 is commented out with `//`; the reference parses
// with nlohmann's ignore_comments at src/testbed.cu:267).
//
// Also implements RFC 7386 merge-patch, which the reference applies for "parent" config
// inheritance (src/testbed.cu:86-97).
#pragma once

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

namespace ngpb {

class Json {
public:
	enum class Type { Null, Bool, Number, String, Array, Object, Binary };

	Json() : m_type(Type::Null) {}
	Json(bool b) : m_type(Type::Bool), m_bool(b) {}
	Json(double d) : m_type(Type::Number), m_num(d) {}
	Json(float d) : m_type(Type::Number), m_num(d) {}
	Json(int d) : m_type(Type::Number), m_num(d), m_is_int(true) {}
	Json(uint32_t d) : m_type(Type::Number), m_num(d), m_is_int(true) {}
	Json(uint64_t d) : m_type(Type::Number), m_num((double)d), m_is_int(true) {}
	Json(int64_t d) : m_type(Type::Number), m_num((double)d), m_is_int(true) {}
	Json(const char* s) : m_type(Type::String), m_str(s) {}
	Json(const std::string& s) : m_type(Type::String), m_str(s) {}

	static Json object() { Json j; j.m_type = Type::Object; return j; }
	static Json array() { Json j; j.m_type = Type::Array; return j; }
	// nlohmann::json::binary_t (msgpack `bin`): shared, so copying a config does not copy its blobs
	static Json binary(std::vector<uint8_t> bytes) { Json j; j.m_type = Type::Binary; j.m_bin = std::make_shared<std::vector<uint8_t>>(std::move(bytes)); return j; }
	static Json binary(const void* data, size_t n) { return binary(std::vector<uint8_t>((const uint8_t*)data, (const uint8_t*)data + n)); }

	Type type() const { return m_type; }
	bool is_null() const { return m_type == Type::Null; }
	bool is_object() const { return m_type == Type::Object; }
	bool is_array() const { return m_type == Type::Array; }
	bool is_number() const { return m_type == Type::Number; }
	bool is_string() const { return m_type == Type::String; }
	bool is_bool() const { return m_type == Type::Bool; }
	bool is_binary() const { return m_type == Type::Binary; }
	// number written without a fraction (an int in the source, or a JSON literal without '.', 'e'): msgpack keeps the distinction
	bool is_integer() const { return m_type == Type::Number && m_is_int; }
	const std::vector<uint8_t>& as_binary() const {
		if (m_type != Type::Binary) throw std::runtime_error{"json: value is not binary"};
		return *m_bin;
	}

	bool contains(const std::string& key) const {
		return m_type == Type::Object && m_obj.count(key) > 0;
	}

	const Json& at(const std::string& key) const {
		auto it = m_obj.find(key);
		if (m_type != Type::Object || it == m_obj.end()) throw std::runtime_error{"json: key '" + key + "' not found"};
		return it->second;
	}
	Json& operator[](const std::string& key) {
		if (m_type == Type::Null) m_type = Type::Object;
		if (m_type != Type::Object) throw std::runtime_error{"json: not an object"};
		return m_obj[key];
	}
	const Json& operator[](const std::string& key) const { return at(key); }
	const Json& operator[](const char* key) const { return at(std::string{key}); }
	Json& operator[](const char* key) { return (*this)[std::string{key}]; }
	const Json& operator[](size_t i) const {
		if (m_type != Type::Array || i >= m_arr.size()) throw std::runtime_error{"json: bad array index"};
		return m_arr[i];
	}
	const Json& operator[](int i) const { return (*this)[(size_t)i]; }   // a literal 0 would otherwise also match the string overload
	const Json& operator[](uint32_t i) const { return (*this)[(size_t)i]; }
	size_t size() const { return m_type == Type::Array ? m_arr.size() : (m_type == Type::Object ? m_obj.size() : 0); }
	void push_back(const Json& v) {
		if (m_type == Type::Null) m_type = Type::Array;
		m_arr.push_back(v);
	}
	void erase(const std::string& key) { m_obj.erase(key); }
	Json& element(size_t i) {
		if (m_type != Type::Array || i >= m_arr.size()) throw std::runtime_error{"json: bad array index"};
		return m_arr[i];
	}
	const std::map<std::string, Json>& items() const { return m_obj; }
	const std::vector<Json>& elements() const { return m_arr; }

	double as_double() const {
		if (m_type == Type::Number) return m_num;
		if (m_type == Type::Bool) return m_bool ? 1.0 : 0.0;
		throw std::runtime_error{"json: value is not a number"};
	}
	float as_float() const { return (float)as_double(); }
	uint32_t as_uint() const { return (uint32_t)as_double(); }
	int as_int() const { return (int)as_double(); }
	bool as_bool() const {
		if (m_type == Type::Bool) return m_bool;
		if (m_type == Type::Number) return m_num != 0;
		throw std::runtime_error{"json: value is not a bool"};
	}
	const std::string& as_string() const {
		if (m_type != Type::String) throw std::runtime_error{"json: value is not a string"};
		return m_str;
	}

	// nlohmann-style `value(key, default)` accessors.
	double value(const std::string& k, double d) const { return contains(k) ? at(k).as_double() : d; }
	float value(const std::string& k, float d) const { return contains(k) ? at(k).as_float() : d; }
	uint32_t value(const std::string& k, uint32_t d) const { return contains(k) ? at(k).as_uint() : d; }
	int value(const std::string& k, int d) const { return contains(k) ? at(k).as_int() : d; }
	bool value(const std::string& k, bool d) const { return contains(k) ? at(k).as_bool() : d; }
	std::string value(const std::string& k, const char* d) const { return contains(k) ? at(k).as_string() : std::string{d}; }
	std::string value(const std::string& k, const std::string& d) const { return contains(k) ? at(k).as_string() : d; }
	Json value(const std::string& k, const Json& d) const { return contains(k) ? at(k) : d; }

	// RFC 7386
	void merge_patch(const Json& patch) {
		if (!patch.is_object()) { *this = patch; return; }
		if (!is_object()) *this = object();
		for (auto& kv : patch.m_obj) {
			if (kv.second.is_null()) m_obj.erase(kv.first);
			else m_obj[kv.first].merge_patch(kv.second);
		}
	}

	std::string dump() const { std::ostringstream o; write(o); return o.str(); }

	static Json parse(const std::string& text) {
		Parser p{text, 0};
		p.skip_ws();
		Json v = p.parse_value();
		p.skip_ws();
		if (p.pos != text.size()) p.fail("trailing characters");
		return v;
	}

private:
	void write(std::ostringstream& o) const {
		switch (m_type) {
			case Type::Null: o << "null"; break;
			case Type::Bool: o << (m_bool ? "true" : "false"); break;
			case Type::Number:
				if (std::floor(m_num) == m_num && std::fabs(m_num) < 1e15) o << (long long)m_num;
				else if (!std::isfinite(m_num)) o << "null";   // as nlohmann's dump does
				else {
					// the shortest text that reads back as the same double (what nlohmann's dump prints): 0.1 stays "0.1",
					// a float widened to double keeps all its digits
					char buf[40];
					for (int digits = 15; digits <= 17; ++digits) {
						snprintf(buf, sizeof(buf), "%.*g", digits, m_num);
						if (strtod(buf, nullptr) == m_num) break;
					}
					o << buf;
				}
				break;
			case Type::String: write_string(o, m_str); break;
			case Type::Binary: { // nlohmann's text form of a binary
				o << "{\"bytes\":[";
				for (size_t i = 0; i < m_bin->size(); ++i) { if (i) o << ","; o << (int)(*m_bin)[i]; }
				o << "],\"subtype\":null}";
			} break;
			case Type::Array: {
				o << "[";
				for (size_t i = 0; i < m_arr.size(); ++i) { if (i) o << ","; m_arr[i].write(o); }
				o << "]";
			} break;
			case Type::Object: {
				o << "{";
				bool first = true;
				for (auto& kv : m_obj) { if (!first) o << ","; first = false; write_string(o, kv.first); o << ":"; kv.second.write(o); }
				o << "}";
			} break;
		}
	}
	static void write_string(std::ostringstream& o, const std::string& s) {
		o << '"';
		for (char c : s) {
			switch (c) {
				case '"': o << "\\\""; break;
				case '\\': o << "\\\\"; break;
				case '\n': o << "\\n"; break;
				case '\t': o << "\\t"; break;
				case '\r': o << "\\r"; break;
				case '\b': o << "\\b"; break;
				case '\f': o << "\\f"; break;
				default:
					if ((unsigned char)c < 0x20) { char esc[8]; snprintf(esc, sizeof(esc), "\\u%04x", (unsigned)(unsigned char)c); o << esc; }
					else o << c;
			}
		}
		o << '"';
	}

	struct Parser {
		const std::string& s;
		size_t pos;
		[[noreturn]] void fail(const std::string& why) const {
			throw std::runtime_error{"json parse error at offset " + std::to_string(pos) + ": " + why};
		}
		void skip_ws() {
			for (;;) {
				while (pos < s.size() && (s[pos] == ' ' || s[pos] == '\t' || s[pos] == '\n' || s[pos] == '\r')) ++pos;
				if (pos + 1 < s.size() && s[pos] == '/' && s[pos + 1] == '/') {
					while (pos < s.size() && s[pos] != '\n') ++pos;
				} else if (pos + 1 < s.size() && s[pos] == '/' && s[pos + 1] == '*') {
					pos += 2;
					while (pos + 1 < s.size() && !(s[pos] == '*' && s[pos + 1] == '/')) ++pos;
					pos += 2;
				} else break;
			}
		}
		// containers recurse: a file of 100 000 opening brackets must end in an error, not in a stack overflow
		int depth = 0;
		struct Nesting {
			Parser& p;
			explicit Nesting(Parser& parser) : p(parser) { if (++p.depth > 512) p.fail("nesting deeper than 512 levels"); }
			~Nesting() { --p.depth; }
		};
		Json parse_value() {
			if (pos >= s.size()) fail("unexpected end");
			char c = s[pos];
			if (c == '{') { Nesting n{*this}; return parse_object(); }
			if (c == '[') { Nesting n{*this}; return parse_array(); }
			if (c == '"') return Json{parse_string()};
			if (s.compare(pos, 4, "true") == 0) { pos += 4; return Json{true}; }
			if (s.compare(pos, 5, "false") == 0) { pos += 5; return Json{false}; }
			if (s.compare(pos, 4, "null") == 0) { pos += 4; return Json{}; }
			return parse_number();
		}
		Json parse_number() {
			size_t start = pos;
			while (pos < s.size() && (isdigit((unsigned char)s[pos]) || s[pos] == '-' || s[pos] == '+' || s[pos] == '.' || s[pos] == 'e' || s[pos] == 'E')) ++pos;
			if (start == pos) fail("unexpected character");
			const std::string lit = s.substr(start, pos - start);
			char* end = nullptr;
			const double d = strtod(lit.c_str(), &end);   // subnormals read as they are; nlohmann accepts them too
			if (end != lit.c_str() + lit.size()) fail("bad number");
			if (!std::isfinite(d)) fail("number out of range");
			Json v{d};
			v.m_is_int = lit.find_first_of(".eE") == std::string::npos;
			return v;
		}
		unsigned parse_hex4() {
			if (pos + 4 >= s.size()) fail("bad \\u escape");
			unsigned cp = 0;
			for (int k = 1; k <= 4; ++k) {
				const char c = s[pos + k];
				cp = cp * 16 + (c >= '0' && c <= '9' ? c - '0' : c >= 'a' && c <= 'f' ? c - 'a' + 10 : c >= 'A' && c <= 'F' ? c - 'A' + 10 : (fail("bad \\u escape"), 0));
			}
			pos += 4;
			return cp;
		}
		std::string parse_string() {
			++pos; // opening quote
			std::string out;
			while (pos < s.size() && s[pos] != '"') {
				if (s[pos] == '\\') {
					++pos;
					if (pos >= s.size()) fail("bad escape");
					switch (s[pos]) {
						case 'n': out += '\n'; break;
						case 't': out += '\t'; break;
						case 'r': out += '\r'; break;
						case 'b': out += '\b'; break;
						case 'f': out += '\f'; break;
						case 'u': { // \uXXXX, surrogate pairs included, stored as UTF-8 (file names written by json.dumps arrive this way)
							unsigned cp = parse_hex4();
							if (cp >= 0xD800 && cp < 0xDC00 && pos + 2 < s.size() && s[pos + 1] == '\\' && s[pos + 2] == 'u') {
								pos += 2;
								const unsigned lo = parse_hex4();
								if (lo < 0xDC00 || lo > 0xDFFF) fail("bad surrogate pair");
								cp = 0x10000 + ((cp - 0xD800) << 10) + (lo - 0xDC00);
							}
							if (cp >= 0xD800 && cp <= 0xDFFF) fail("bad surrogate pair");
							if (cp < 0x80) out += (char)cp;
							else if (cp < 0x800) { out += (char)(0xC0 | (cp >> 6)); out += (char)(0x80 | (cp & 0x3F)); }
							else if (cp < 0x10000) { out += (char)(0xE0 | (cp >> 12)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
							else { out += (char)(0xF0 | (cp >> 18)); out += (char)(0x80 | ((cp >> 12) & 0x3F)); out += (char)(0x80 | ((cp >> 6) & 0x3F)); out += (char)(0x80 | (cp & 0x3F)); }
						} break;
						default: out += s[pos];
					}
					++pos;
				} else out += s[pos++];
			}
			if (pos >= s.size()) fail("unterminated string");
			++pos;
			return out;
		}
		Json parse_array() {
			++pos;
			Json a = Json::array();
			skip_ws();
			if (pos < s.size() && s[pos] == ']') { ++pos; return a; }
			for (;;) {
				skip_ws();
				a.push_back(parse_value());
				skip_ws();
				if (pos >= s.size()) fail("unterminated array");
				if (s[pos] == ',') { ++pos; continue; }
				if (s[pos] == ']') { ++pos; return a; }
				fail("expected , or ]");
			}
		}
		Json parse_object() {
			++pos;
			Json o = Json::object();
			skip_ws();
			if (pos < s.size() && s[pos] == '}') { ++pos; return o; }
			for (;;) {
				skip_ws();
				if (pos >= s.size() || s[pos] != '"') fail("expected string key");
				std::string k = parse_string();
				skip_ws();
				if (pos >= s.size() || s[pos] != ':') fail("expected :");
				++pos;
				skip_ws();
				o[k] = parse_value();
				skip_ws();
				if (pos >= s.size()) fail("unterminated object");
				if (s[pos] == ',') { ++pos; continue; }
				if (s[pos] == '}') { ++pos; return o; }
				fail("expected , or }");
			}
		}
	};

	Type m_type;
	bool m_bool = false;
	double m_num = 0;
	bool m_is_int = false;
	std::shared_ptr<std::vector<uint8_t>> m_bin;
	std::string m_str;
	std::vector<Json> m_arr;
	std::map<std::string, Json> m_obj;
};

inline bool equals_case_insensitive(const std::string& a, const std::string& b) {
	if (a.size() != b.size()) return false;
	for (size_t i = 0; i < a.size(); ++i) if (tolower((unsigned char)a[i]) != tolower((unsigned char)b[i])) return false;
	return true;
}

} // namespace ngpb
