// The reference's own JSON library (nlohmann json as vendored under tiny-cuda-nn/dependencies/json) compiled for the
// host from where it lies under /root/reference: what Testbed::save_snapshot's json::to_msgpack and
// merge_parent_network_config's json::merge_patch (src/testbed.cu:86-97, 4826-4836) actually produce. TEST
// INFRASTRUCTURE: used by tests/golden/make_ref_json_vectors.py to write the fixtures that pin csrc/msgpack.h and
// csrc/json.h; nothing of the product links it. Built by oracle/build_ref.sh into oracle/_ref/libref_json.so.
#include <json/json.hpp>

#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

using json = nlohmann::json;

// strings of the form "bin:<hex>" stand for json::binary_t values (text JSON has no binary literal)
static void restore_binaries(json& j) {
	if (j.is_string()) {
		const std::string& s = j.get_ref<const std::string&>();
		if (s.rfind("bin:", 0) == 0) {
			std::vector<uint8_t> bytes;
			for (size_t i = 4; i + 1 < s.size(); i += 2) bytes.push_back((uint8_t)std::stoul(s.substr(i, 2), nullptr, 16));
			j = json::binary(std::move(bytes));
		}
	} else if (j.is_structured()) {
		for (auto& e : j) restore_binaries(e);
	}
}

extern "C" size_t ref_json_to_msgpack(const char* text, uint8_t* out, size_t cap) {
	json j = json::parse(text, nullptr, true, /*ignore_comments*/ true);
	restore_binaries(j);
	const std::vector<uint8_t> bytes = json::to_msgpack(j);
	if (out && cap >= bytes.size()) memcpy(out, bytes.data(), bytes.size());
	return bytes.size();
}

// parent.merge_patch(child) as merge_parent_network_config applies it; the result as compact JSON text
extern "C" size_t ref_json_merge_patch(const char* parent_text, const char* child_text, char* out, size_t cap) {
	json parent = json::parse(parent_text, nullptr, true, true);
	const json child = json::parse(child_text, nullptr, true, true);
	parent.merge_patch(child);
	const std::string s = parent.dump();
	if (out && cap > s.size()) memcpy(out, s.c_str(), s.size() + 1);
	return s.size() + 1;
}

// msgpack -> json -> msgpack through the reference's library (what its load_snapshot + save_snapshot would do to a container)
extern "C" size_t ref_msgpack_roundtrip(const uint8_t* in, size_t n, uint8_t* out, size_t cap) {
	const json j = json::from_msgpack(in, in + n);
	const std::vector<uint8_t> bytes = json::to_msgpack(j);
	if (out && cap >= bytes.size()) memcpy(out, bytes.data(), bytes.size());
	return bytes.size();
}

// The frame order of the reference's loader (src/nerf_loader.cu:347-349): SI::natural::compare from the vendored
// dependencies/NaturalSort, as it is. 1 when a sorts before b.
#include <natural_sort.hpp>
extern "C" int ref_natural_less(const char* a, const char* b) {
	return SI::natural::compare<std::string>(std::string{a}, std::string{b}) ? 1 : 0;
}
