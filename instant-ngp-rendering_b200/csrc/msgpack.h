// msgpack <-> Json, the subset nlohmann::json::to_msgpack / from_msgpack produce and accept for the reference's
// snapshot files (src/testbed.cu:4826-4836, :246-272): nil, bool, integers of every width, float32 / float64, str,
// bin (json::binary_t without subtype), array, map with string keys. Host only.
//
// Writer choices follow nlohmann (the files are then byte-compatible where the values are): the smallest integer
// encoding that holds the value, float32 when the double converts to float and back unchanged, else float64.
#pragma once

#include "json.h"

#include <cstring>
#include <limits>

namespace ngpb {

class MsgpackWriter {
public:
	std::vector<uint8_t> out;

	void write(const Json& j) {
		switch (j.type()) {
			case Json::Type::Null: out.push_back(0xc0); break;
			case Json::Type::Bool: out.push_back(j.as_bool() ? 0xc3 : 0xc2); break;
			case Json::Type::Number:
				if (j.is_integer()) write_int(j.as_double()); else write_float(j.as_double());
				break;
			case Json::Type::String: {
				const std::string& s = j.as_string();
				write_str(s);
			} break;
			case Json::Type::Binary: {
				const std::vector<uint8_t>& b = j.as_binary();
				if (b.size() <= 0xFF) { out.push_back(0xc4); be<uint8_t>((uint8_t)b.size()); }
				else if (b.size() <= 0xFFFF) { out.push_back(0xc5); be<uint16_t>((uint16_t)b.size()); }
				else if (b.size() <= 0xFFFFFFFFull) { out.push_back(0xc6); be<uint32_t>((uint32_t)b.size()); }
				else throw std::runtime_error{"msgpack: binary larger than 4 GiB"};
				out.insert(out.end(), b.begin(), b.end());
			} break;
			case Json::Type::Array: {
				const size_t n = j.size();
				if (n <= 15) out.push_back((uint8_t)(0x90 | n));
				else if (n <= 0xFFFF) { out.push_back(0xdc); be<uint16_t>((uint16_t)n); }
				else { out.push_back(0xdd); be<uint32_t>((uint32_t)n); }
				for (const Json& e : j.elements()) write(e);
			} break;
			case Json::Type::Object: {
				const size_t n = j.size();
				if (n <= 15) out.push_back((uint8_t)(0x80 | n));
				else if (n <= 0xFFFF) { out.push_back(0xde); be<uint16_t>((uint16_t)n); }
				else { out.push_back(0xdf); be<uint32_t>((uint32_t)n); }
				for (const auto& kv : j.items()) { write_str(kv.first); write(kv.second); }
			} break;
		}
	}

private:
	template <typename T> void be(T v) {
		for (int i = (int)sizeof(T) - 1; i >= 0; --i) out.push_back((uint8_t)((uint64_t)v >> (8 * i)));
	}
	void write_str(const std::string& s) {
		const size_t n = s.size();
		if (n <= 31) out.push_back((uint8_t)(0xa0 | n));
		else if (n <= 0xFF) { out.push_back(0xd9); be<uint8_t>((uint8_t)n); }
		else if (n <= 0xFFFF) { out.push_back(0xda); be<uint16_t>((uint16_t)n); }
		else { out.push_back(0xdb); be<uint32_t>((uint32_t)n); }
		out.insert(out.end(), s.begin(), s.end());
	}
	void write_int(double d) {
		if (d >= 0) {
			const uint64_t v = (uint64_t)d;
			if (v <= 127) out.push_back((uint8_t)v);
			else if (v <= 0xFF) { out.push_back(0xcc); be<uint8_t>((uint8_t)v); }
			else if (v <= 0xFFFF) { out.push_back(0xcd); be<uint16_t>((uint16_t)v); }
			else if (v <= 0xFFFFFFFFull) { out.push_back(0xce); be<uint32_t>((uint32_t)v); }
			else { out.push_back(0xcf); be<uint64_t>(v); }
		} else {
			const int64_t v = (int64_t)d;
			if (v >= -32) out.push_back((uint8_t)(int8_t)v);
			else if (v >= std::numeric_limits<int8_t>::min()) { out.push_back(0xd0); be<uint8_t>((uint8_t)(int8_t)v); }
			else if (v >= std::numeric_limits<int16_t>::min()) { out.push_back(0xd1); be<uint16_t>((uint16_t)(int16_t)v); }
			else if (v >= std::numeric_limits<int32_t>::min()) { out.push_back(0xd2); be<uint32_t>((uint32_t)(int32_t)v); }
			else { out.push_back(0xd3); be<uint64_t>((uint64_t)v); }
		}
	}
	void write_float(double d) {
		const float f = (float)d;
		if (std::fabs(d) <= (double)std::numeric_limits<float>::max() && (double)f == d) {
			uint32_t u; memcpy(&u, &f, 4);
			out.push_back(0xca); be<uint32_t>(u);
		} else {
			uint64_t u; memcpy(&u, &d, 8);
			out.push_back(0xcb); be<uint64_t>(u);
		}
	}
};

class MsgpackReader {
public:
	MsgpackReader(const uint8_t* data, size_t size) : m_p(data), m_end(data + size) {}

	Json read() {
		struct Nesting {   // arrays and maps recurse: bound the depth as the JSON parser does
			int& d;
			explicit Nesting(int& depth) : d(depth) { if (++d > 512) throw std::runtime_error{"msgpack: nesting deeper than 512 levels"}; }
			~Nesting() { --d; }
		} nesting{m_depth};
		const uint8_t t = byte();
		if (t <= 0x7f) return Json{(uint32_t)t};
		if (t >= 0xe0) return Json{(int)(int8_t)t};
		if ((t & 0xe0) == 0xa0) return Json{str(t & 0x1f)};
		if ((t & 0xf0) == 0x90) return array(t & 0x0f);
		if ((t & 0xf0) == 0x80) return map(t & 0x0f);
		switch (t) {
			case 0xc0: return Json{};
			case 0xc2: return Json{false};
			case 0xc3: return Json{true};
			case 0xc4: return bin(be<uint8_t>());
			case 0xc5: return bin(be<uint16_t>());
			case 0xc6: return bin(be<uint32_t>());
			case 0xca: { const uint32_t u = be<uint32_t>(); float f; memcpy(&f, &u, 4); return Json{(double)f}; }
			case 0xcb: { const uint64_t u = be<uint64_t>(); double d; memcpy(&d, &u, 8); return Json{d}; }
			case 0xcc: return Json{(uint32_t)be<uint8_t>()};
			case 0xcd: return Json{(uint32_t)be<uint16_t>()};
			case 0xce: return Json{be<uint32_t>()};
			case 0xcf: return Json{be<uint64_t>()};
			case 0xd0: return Json{(int)(int8_t)be<uint8_t>()};
			case 0xd1: return Json{(int)(int16_t)be<uint16_t>()};
			case 0xd2: return Json{(int)(int32_t)be<uint32_t>()};
			case 0xd3: return Json{(int64_t)be<uint64_t>()};
			case 0xd9: return Json{str(be<uint8_t>())};
			case 0xda: return Json{str(be<uint16_t>())};
			case 0xdb: return Json{str(be<uint32_t>())};
			case 0xdc: return array(be<uint16_t>());
			case 0xdd: return array(be<uint32_t>());
			case 0xde: return map(be<uint16_t>());
			case 0xdf: return map(be<uint32_t>());
			default: throw std::runtime_error{"msgpack: unsupported type byte " + std::to_string((int)t) + " (ext types are not part of a snapshot)"};
		}
	}
	bool at_end() const { return m_p == m_end; }

private:
	void need(size_t n) const {
		if ((size_t)(m_end - m_p) < n) throw std::runtime_error{"msgpack: unexpected end of input"};
	}
	uint8_t byte() { need(1); return *m_p++; }
	template <typename T> T be() {
		need(sizeof(T));
		uint64_t v = 0;
		for (size_t i = 0; i < sizeof(T); ++i) v = (v << 8) | *m_p++;
		return (T)v;
	}
	std::string str(size_t n) {
		need(n);
		std::string s((const char*)m_p, n);
		m_p += n;
		return s;
	}
	Json bin(size_t n) {
		need(n);
		Json j = Json::binary(m_p, n);
		m_p += n;
		return j;
	}
	Json array(size_t n) {
		Json a = Json::array();
		for (size_t i = 0; i < n; ++i) a.push_back(read());
		return a;
	}
	Json map(size_t n) {
		Json o = Json::object();
		for (size_t i = 0; i < n; ++i) {
			Json k = read();
			if (!k.is_string()) throw std::runtime_error{"msgpack: map key is not a string"};
			o[k.as_string()] = read();
		}
		return o;
	}
	const uint8_t* m_p;
	const uint8_t* m_end;
	int m_depth = 0;
};

inline std::vector<uint8_t> to_msgpack(const Json& j) {
	MsgpackWriter w;
	w.write(j);
	return std::move(w.out);
}
inline Json from_msgpack(const uint8_t* data, size_t size) {
	MsgpackReader r{data, size};
	Json j = r.read();
	if (!r.at_end()) throw std::runtime_error{"msgpack: trailing bytes after the document"};
	return j;
}

} // namespace ngpb
