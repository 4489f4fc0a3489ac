// Host-side file helpers shared by host_io.cu (config files, containers), host_snapshot.cu and host_dataset.cu.
#pragma once

#include "json.h"

#include <string>
#include <vector>

namespace ngpb {

bool file_exists(const std::string& path);
bool is_directory(const std::string& path);
std::string path_dirname(const std::string& path);
std::string path_basename(const std::string& path);
std::string path_extension(const std::string& path);   // lower case, without the dot
std::string path_join(const std::string& dir, const std::string& rel);
std::vector<uint8_t> read_file(const std::string& path);

// load_network_config (src/testbed.cu:246-272): *.json (comments, "parent"), *.msgpack, *.ingp (zlib + msgpack)
Json load_network_config(const std::string& path);
// json::to_msgpack into the file; *.ingp gets the zlib stream (level 0 when not `compress`), src/testbed.cu:4826-4836
void write_snapshot_container(const std::string& path, const Json& config, bool compress);

// Decoded image: 8-bit RGBA, rows top to bottom.
struct HostImage {
	int width = 0, height = 0;
	std::vector<uint8_t> rgba;
};
// Decoded HDR image: float RGBA, rows top to bottom.
struct HostImageF {
	int width = 0, height = 0;
	std::vector<float> rgba;
};
// OpenEXR, single-part scanline or tiled, HALF / FLOAT channels, compression NONE / RLE / ZIPS / ZIP / PIZ (host_exr.cu).
// The reference reads EXR files through two doors with different channel rules:
//   Image    tinyexr's LoadEXR behind load_exr (image mode, src/tinyexr_wrapper.cu:125-151): a one-channel file fills R, G, B and A
//            with that channel; otherwise R, G, B must exist ("R channel not found") and a missing A is 1
//   Dataset  load_exr_to_gpu (NeRF frames, :153-258): R / G / B / A by name, absent colour channels are 0, absent A is 1,
//            mixed HALF / FLOAT channels are refused
enum class ExrChannelRule { Image, Dataset };
HostImageF decode_exr(const std::vector<uint8_t>& file, const std::string& name_for_errors, ExrChannelRule rule);

// PNG (8 / 16 bit, grey / grey-alpha / RGB / RGBA / palette, plain or Adam7-interlaced) decoded on the host with zlib.
HostImage decode_png(const std::vector<uint8_t>& file, const std::string& name_for_errors);
// Any 8-bit format of the loader, picked from the file's content as stb_image does: JPEG (host_jpeg.cu), PNG, BMP, GIF, PSD, PIC, TGA, binary PNM (host_image.cu).
HostImage decode_image_rgba8(const std::vector<uint8_t>& file, const std::string& name_for_errors);
bool is_jpeg_file(const std::vector<uint8_t>& file);
// Radiance HDR (RGBE) as floats, alpha 1: what stbi_loadf(..., 4) returns (the image mode's reader for *.hdr files).
HostImageF decode_hdr(const std::vector<uint8_t>& file, const std::string& name_for_errors);
bool is_hdr_file(const std::vector<uint8_t>& file);   // stbi_is_hdr
// JPEG on the host with stb_image's arithmetic after entropy decoding, so the texels are the reference loader's (host_jpeg.cu).
HostImage decode_jpeg(const std::vector<uint8_t>& file, const std::string& name_for_errors);

} // namespace ngpb
