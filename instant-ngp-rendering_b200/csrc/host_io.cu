// File-level entry points of the Testbed in host C++, as in the reference: network config files with "parent"
// inheritance (src/testbed.cu:86-97, :246-308) and snapshots (src/testbed.cu:4775-4972; Trainer::serialize /
// deserialize, tcnn trainer.h:275-314; optimizer chains adam.h:278-299, ema.h:190-206, exponential_decay.h:136-148;
// key names of include/neural-graphics-primitives/json_binding.h, matrices as lists of ROWS as tcnn's vec_json.h).
// The container is json::to_msgpack, wrapped in a zlib stream when the file is called *.ingp (zstr).
// The Python mirror (snapshot.py) is a second implementation of the same format; tests read each one's files with
// the other.
#include "testbed.h"
#include "host_io.h"
#include "msgpack.h"

#include <zlib.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <sys/stat.h>

namespace ngpb {

constexpr int kSnapshotFormatVersion = 1;   // src/testbed.cu:4773

// ---- files -------------------------------------------------------------------------------------------------
bool file_exists(const std::string& path) {
	struct stat st;
	return !path.empty() && stat(path.c_str(), &st) == 0;
}
bool is_directory(const std::string& path) {
	struct stat st;
	return !path.empty() && stat(path.c_str(), &st) == 0 && S_ISDIR(st.st_mode);
}
std::string path_dirname(const std::string& path) {
	const size_t k = path.find_last_of('/');
	return k == std::string::npos ? std::string{"."} : (k == 0 ? std::string{"/"} : path.substr(0, k));
}
std::string path_basename(const std::string& path) {
	const size_t k = path.find_last_of('/');
	return k == std::string::npos ? path : path.substr(k + 1);
}
std::string path_extension(const std::string& path) {   // lower case, without the dot
	const std::string base = path_basename(path);
	const size_t k = base.find_last_of('.');
	if (k == std::string::npos || k == 0) return "";
	std::string e = base.substr(k + 1);
	for (char& c : e) c = (char)tolower((unsigned char)c);
	return e;
}
std::string path_join(const std::string& dir, const std::string& rel) {
	if (!rel.empty() && rel[0] == '/') return rel;
	if (dir.empty() || dir == ".") return rel;
	return dir.back() == '/' ? dir + rel : dir + "/" + rel;
}
std::vector<uint8_t> read_file(const std::string& path) {
	std::ifstream f{path, std::ios::binary | std::ios::ate};
	if (!f) throw std::runtime_error{"Could not open file '" + path + "'."};
	const std::streamsize n = f.tellg();
	f.seekg(0);
	std::vector<uint8_t> data((size_t)std::max<std::streamsize>(n, 0));
	if (n > 0 && !f.read((char*)data.data(), n)) throw std::runtime_error{"Could not read file '" + path + "'."};
	return data;
}
static void write_file(const std::string& path, const uint8_t* data, size_t n) {
	std::ofstream f{path, std::ios::binary};
	if (!f || !f.write((const char*)data, (std::streamsize)n)) throw std::runtime_error{"Could not write file '" + path + "'."};
}

// ---- zlib --------------------------------------------------------------------------------------------------
static std::vector<uint8_t> zlib_deflate(const std::vector<uint8_t>& in, int level) {
	uLongf cap = compressBound((uLong)in.size());
	std::vector<uint8_t> out(cap);
	if (compress2(out.data(), &cap, in.data(), (uLong)in.size(), level) != Z_OK) throw std::runtime_error{"zlib: compress2 failed"};
	out.resize(cap);
	return out;
}
// zlib or gzip framing (zstr accepts both); the caller has checked the header
static std::vector<uint8_t> zlib_inflate(const std::vector<uint8_t>& in) {
	z_stream zs{};
	if (inflateInit2(&zs, 15 + 32) != Z_OK) throw std::runtime_error{"zlib: inflateInit2 failed"};
	std::vector<uint8_t> out(std::max<size_t>(in.size() * 2, 1u << 16));
	zs.next_in = const_cast<Bytef*>(in.data());
	zs.avail_in = (uInt)in.size();
	size_t produced = 0;
	for (;;) {
		if (produced == out.size()) out.resize(out.size() * 2);
		zs.next_out = out.data() + produced;
		zs.avail_out = (uInt)std::min<size_t>(out.size() - produced, 1u << 30);
		const size_t before = zs.avail_out;
		const int rc = inflate(&zs, Z_NO_FLUSH);
		produced += before - zs.avail_out;
		if (rc == Z_STREAM_END) break;
		if (rc != Z_OK && rc != Z_BUF_ERROR) { inflateEnd(&zs); throw std::runtime_error{"zlib: the stream is corrupt"}; }
		if (rc == Z_BUF_ERROR && zs.avail_in == 0) { inflateEnd(&zs); throw std::runtime_error{"zlib: the stream is truncated"}; }
	}
	inflateEnd(&zs);
	out.resize(produced);
	return out;
}

// ---- network config files -------------------------------------------------------------------------------------
// load_network_config (src/testbed.cu:246-272) + merge_parent_network_config (:86-97): *.json is text with comments and
// "parent" inheritance (relative to the child's directory), *.msgpack is msgpack, *.ingp is zlib + msgpack, anything else
// reads as an empty config.
Json load_network_config(const std::string& path) {
	const std::string ext = path_extension(path);
	const bool is_snapshot = ext == "msgpack" || ext == "ingp";
	if (!file_exists(path)) throw std::runtime_error{std::string{"Network "} + (is_snapshot ? "snapshot" : "config") + " '" + path + "' does not exist."};
	if (ext == "json") {
		const std::vector<uint8_t> text = read_file(path);
		Json cfg = Json::parse(std::string{text.begin(), text.end()});
		if (cfg.contains("parent")) {
			Json parent = load_network_config(path_join(path_dirname(path), cfg["parent"].as_string()));
			cfg.erase("parent");
			parent.merge_patch(cfg);
			return parent;
		}
		return cfg;
	}
	if (is_snapshot) {
		std::vector<uint8_t> blob = read_file(path);
		// zstr passes a stream without a zlib / gzip header through unchanged
		if (ext == "ingp" && blob.size() >= 2 && (blob[0] == 0x78 || (blob[0] == 0x1f && blob[1] == 0x8b))) blob = zlib_inflate(blob);
		return from_msgpack(blob.data(), blob.size());
	}
	return Json::object();
}

void Testbed::reload_network_from_file(const std::string& path) {
	// :274-308: a snapshot file given as a config carries its network config at the top level; "snapshot" is dropped
	Json cfg = load_network_config(path);
	if (cfg.contains("snapshot")) cfg.erase("snapshot");
	reload_network_from_json(cfg);
}

// Testbed::load_file (src/testbed.cu:318-391): snapshot, network config, or training data by what the file is.
void Testbed::load_file(const std::string& path) {
	const std::string ext = path_extension(path);
	if (!file_exists(path)) throw std::runtime_error{"File '" + path + "' does not exist."};
	if (ext == "ingp" || ext == "msgpack") { load_snapshot(path); return; }
	if (ext == "json") {
		const std::vector<uint8_t> text = read_file(path);
		const Json doc = Json::parse(std::string{text.begin(), text.end()});
		if (doc.contains("snapshot")) throw std::runtime_error{"ngp-b200: snapshots in JSON form are not supported; use .ingp / .msgpack"};
		if (doc.contains("path") && !doc.contains("frames")) throw std::runtime_error{"ngp-b200: camera paths are not part of the path (DESIGN.md section 7)"};
		if (doc.contains("parent") || doc.contains("network") || doc.contains("encoding") || doc.contains("loss") || doc.contains("optimizer")) {
			reload_network_from_file(path);
			return;
		}
	}
	load_training_data(path);
}

void write_snapshot_container(const std::string& path, const Json& config, bool compress) {
	std::vector<uint8_t> blob = to_msgpack(config);
	if (path_extension(path) == "ingp") blob = zlib_deflate(blob, compress ? Z_DEFAULT_COMPRESSION : 0);
	write_file(path, blob.data(), blob.size());
}

} // namespace ngpb

// ---- host-only C entry points (no device needed): what a C++ host uses to look at files, and what the CPU tests call ----
#include "../../include/ngp_b200.h"

namespace {
size_t copy_text(const std::string& s, char* buf, size_t cap) {
	if (buf && cap > 0) {
		const size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
		memcpy(buf, s.data(), n);
		buf[n] = 0;
	}
	return s.size() + 1;
}
} // namespace

extern "C" size_t ngpb_host_load_network_config(const char* path, char* buf, size_t cap) {
	try {
		if (!path) throw std::runtime_error{"ngpb_host_load_network_config: null argument"};
		return copy_text(ngpb::load_network_config(path).dump(), buf, cap);
	} catch (const std::exception& e) { ngpb::set_last_error(e.what()); return 0; }
}

extern "C" size_t ngpb_host_msgpack_reencode(const void* in, size_t n, void* out, size_t cap) {
	try {
		if (!in) throw std::runtime_error{"ngpb_host_msgpack_reencode: null argument"};
		const std::vector<uint8_t> blob = ngpb::to_msgpack(ngpb::from_msgpack((const uint8_t*)in, n));
		if (out && cap >= blob.size()) memcpy(out, blob.data(), blob.size());
		return blob.size();
	} catch (const std::exception& e) { ngpb::set_last_error(e.what()); return 0; }
}

extern "C" int ngpb_host_decode_png(const char* path, void* rgba, size_t cap, int* width, int* height) {
	try {
		if (!path || !width || !height) throw std::runtime_error{"ngpb_host_decode_png: null argument"};
		const ngpb::HostImage img = ngpb::decode_png(ngpb::read_file(path), path);
		*width = img.width; *height = img.height;
		if (rgba && cap >= img.rgba.size()) memcpy(rgba, img.rgba.data(), img.rgba.size());
		return NGPB_OK;
	} catch (const std::exception& e) { ngpb::set_last_error(e.what()); return NGPB_ERROR; }
}

extern "C" int ngpb_host_decode_image(const char* path, void* rgba, size_t cap, int* width, int* height) {
	try {
		if (!path || !width || !height) throw std::runtime_error{"ngpb_host_decode_image: null argument"};
		const ngpb::HostImage img = ngpb::decode_image_rgba8(ngpb::read_file(path), path);
		*width = img.width; *height = img.height;
		if (rgba && cap >= img.rgba.size()) memcpy(rgba, img.rgba.data(), img.rgba.size());
		return NGPB_OK;
	} catch (const std::exception& e) { ngpb::set_last_error(e.what()); return NGPB_ERROR; }
}

extern "C" int ngpb_host_decode_exr(const char* path, float* rgba, size_t cap_floats, int* width, int* height) {
	try {
		if (!path || !width || !height) throw std::runtime_error{"ngpb_host_decode_exr: null argument"};
		const ngpb::HostImageF img = ngpb::decode_exr(ngpb::read_file(path), path, ngpb::ExrChannelRule::Image);
		*width = img.width; *height = img.height;
		if (rgba && cap_floats >= img.rgba.size()) memcpy(rgba, img.rgba.data(), img.rgba.size() * sizeof(float));
		return NGPB_OK;
	} catch (const std::exception& e) { ngpb::set_last_error(e.what()); return NGPB_ERROR; }
}

extern "C" int ngpb_host_decode_hdr(const char* path, float* rgba, size_t cap_floats, int* width, int* height) {
	try {
		if (!path || !width || !height) throw std::runtime_error{"ngpb_host_decode_hdr: null argument"};
		const ngpb::HostImageF img = ngpb::decode_hdr(ngpb::read_file(path), path);
		*width = img.width; *height = img.height;
		if (rgba && cap_floats >= img.rgba.size()) memcpy(rgba, img.rgba.data(), img.rgba.size() * sizeof(float));
		return NGPB_OK;
	} catch (const std::exception& e) { ngpb::set_last_error(e.what()); return NGPB_ERROR; }
}

extern "C" int ngpb_host_decode_exr_frame(const char* path, float* rgba, size_t cap_floats, int* width, int* height) {
	try {
		if (!path || !width || !height) throw std::runtime_error{"ngpb_host_decode_exr_frame: null argument"};
		const ngpb::HostImageF img = ngpb::decode_exr(ngpb::read_file(path), path, ngpb::ExrChannelRule::Dataset);
		*width = img.width; *height = img.height;
		if (rgba && cap_floats >= img.rgba.size()) memcpy(rgba, img.rgba.data(), img.rgba.size() * sizeof(float));
		return NGPB_OK;
	} catch (const std::exception& e) { ngpb::set_last_error(e.what()); return NGPB_ERROR; }
}
