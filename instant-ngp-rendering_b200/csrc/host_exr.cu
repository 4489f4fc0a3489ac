// OpenEXR reader for the loader (the reference reads EXR through tinyexr, src/tinyexr_wrapper.cu, which is not part of
// this image): single-part scanline files, HALF and FLOAT channels, compression NONE / ZIPS / ZIP - what
// data/image/albert.exr (FLOAT RGBA, ZIP) and scripts-written HDR scenes use. PIZ / PXR24 / B44 / DWA, tiled and
// multi-part files are refused with a message. Output: float RGBA, rows top to bottom (a missing alpha is 1, a single
// luminance channel Y is replicated). Host only; zlib does the inflate.
#include "host_io.h"

#include <zlib.h>

#include <cuda_fp16.h>

#include <cstring>
#include <map>
#include <stdexcept>

namespace ngpb {

namespace {
struct ExrChannel {
	std::string name;
	int pixel_type;   // 0 UINT, 1 HALF, 2 FLOAT
};
template <typename T> T rd(const uint8_t* p) { T v; memcpy(&v, p, sizeof(T)); return v; }
} // namespace

HostImageF decode_exr(const std::vector<uint8_t>& file, const std::string& name) {
	auto fail = [&](const std::string& why) -> void { throw std::runtime_error{"Could not load image file '" + name + "': " + why}; };
	if (file.size() < 8 || rd<uint32_t>(file.data()) != 20000630u) fail("not an OpenEXR file");
	const uint32_t version = rd<uint32_t>(file.data() + 4);
	if ((version & 0xFF) != 2) fail("unsupported OpenEXR version");
	if (version & 0x200) fail("tiled files are not supported");
	if (version & 0x1800) fail("deep / multi-part files are not supported");

	std::vector<ExrChannel> channels;
	int compression = -1, xmin = 0, ymin = 0, xmax = -1, ymax = -1;
	size_t pos = 8;
	auto cstr = [&]() {
		const size_t start = pos;
		while (pos < file.size() && file[pos] != 0) ++pos;
		if (pos >= file.size()) fail("truncated header");
		return std::string((const char*)&file[start], pos++ - start);
	};
	for (;;) {
		const std::string attr = cstr();
		if (attr.empty()) break;
		const std::string type = cstr();
		if (pos + 4 > file.size()) fail("truncated header");
		const uint32_t size = rd<uint32_t>(&file[pos]);
		pos += 4;
		if (pos + size > file.size()) fail("truncated header");
		const uint8_t* v = &file[pos];
		if (attr == "channels") {
			size_t p = 0;
			while (p < size && v[p] != 0) {
				ExrChannel ch;
				while (p < size && v[p] != 0) ch.name += (char)v[p++];
				++p;
				if (p + 16 > size) fail("bad channel list");
				ch.pixel_type = rd<int32_t>(v + p);
				if (rd<int32_t>(v + p + 8) != 1 || rd<int32_t>(v + p + 12) != 1) fail("subsampled channels are not supported");
				p += 16;
				channels.push_back(ch);
			}
		} else if (attr == "compression") compression = v[0];
		else if (attr == "dataWindow") { xmin = rd<int32_t>(v); ymin = rd<int32_t>(v + 4); xmax = rd<int32_t>(v + 8); ymax = rd<int32_t>(v + 12); }
		pos += size;
	}
	const int w = xmax - xmin + 1, h = ymax - ymin + 1;
	if (channels.empty() || w <= 0 || h <= 0 || w > (1 << 16) || h > (1 << 16)) fail("missing channels or data window");
	int lines_per_block = 0;
	switch (compression) {
		case 0: case 2: lines_per_block = 1; break;
		case 3: lines_per_block = 16; break;
		default: fail("compression type " + std::to_string(compression) + " is not supported (NONE, ZIPS and ZIP are)");
	}
	size_t bytes_per_line = 0;
	for (const ExrChannel& ch : channels) {
		if (ch.pixel_type != 1 && ch.pixel_type != 2) fail("UINT channels are not supported");
		bytes_per_line += (size_t)w * (ch.pixel_type == 1 ? 2 : 4);
	}
	// where each stored channel goes in the RGBA output (-1: ignored)
	std::vector<int> target(channels.size(), -1);
	bool has_alpha = false, has_rgb = false;
	for (size_t c = 0; c < channels.size(); ++c) {
		const std::string& n = channels[c].name;
		if (n == "R") { target[c] = 0; has_rgb = true; }
		else if (n == "G") { target[c] = 1; has_rgb = true; }
		else if (n == "B") { target[c] = 2; has_rgb = true; }
		else if (n == "A") { target[c] = 3; has_alpha = true; }
	}
	int grey = -1;
	if (!has_rgb) for (size_t c = 0; c < channels.size(); ++c) if (channels[c].name == "Y") grey = (int)c;
	if (!has_rgb && grey < 0) fail("no R, G, B or Y channel");

	const int n_blocks = (h + lines_per_block - 1) / lines_per_block;
	if (pos + (size_t)n_blocks * 8 > file.size()) fail("truncated offset table");
	HostImageF img;
	img.width = w; img.height = h;
	img.rgba.assign((size_t)w * h * 4, 0.0f);
	if (!has_alpha) for (size_t i = 0; i < (size_t)w * h; ++i) img.rgba[i * 4 + 3] = 1.0f;
	std::vector<uint8_t> raw, tmp;
	for (int b = 0; b < n_blocks; ++b) {
		const uint64_t off = rd<uint64_t>(&file[pos + (size_t)b * 8]);
		if (off + 8 > file.size()) fail("bad chunk offset");
		const int y0 = rd<int32_t>(&file[off]) - ymin;
		const uint32_t packed = rd<uint32_t>(&file[off + 4]);
		if (y0 < 0 || y0 >= h || off + 8 + packed > file.size()) fail("bad chunk");
		const int lines = std::min(lines_per_block, h - y0);
		const size_t unpacked = bytes_per_line * lines;
		const uint8_t* src = &file[off + 8];
		raw.resize(unpacked);
		if (compression == 0 || packed == unpacked) {
			if (packed != unpacked) fail("bad uncompressed chunk size");
			memcpy(raw.data(), src, unpacked);
		} else {
			tmp.resize(unpacked);
			uLongf got = (uLongf)unpacked;
			if (uncompress(tmp.data(), &got, src, packed) != Z_OK || got != unpacked) fail("corrupt ZIP chunk");
			// predictor, then the two interleaved halves (ImfZip.cpp)
			for (size_t i = 1; i < unpacked; ++i) tmp[i] = (uint8_t)(tmp[i - 1] + tmp[i] - 128);
			const size_t half = (unpacked + 1) / 2;
			for (size_t i = 0; i < unpacked; ++i) raw[i] = (i & 1) ? tmp[half + i / 2] : tmp[i / 2];
		}
		const uint8_t* p = raw.data();
		for (int l = 0; l < lines; ++l) {
			float* out = &img.rgba[(size_t)(y0 + l) * w * 4];
			for (size_t c = 0; c < channels.size(); ++c) {
				const bool is_half = channels[c].pixel_type == 1;
				const int t = (int)c == grey ? 4 : target[c];
				if (t >= 0) {
					for (int x = 0; x < w; ++x) {
						const float v = is_half ? __half2float(rd<__half>(p + (size_t)x * 2)) : rd<float>(p + (size_t)x * 4);
						if (t == 4) { out[x * 4] = out[x * 4 + 1] = out[x * 4 + 2] = v; }
						else out[x * 4 + t] = v;
					}
				}
				p += (size_t)w * (is_half ? 2 : 4);
			}
		}
	}
	return img;
}

} // namespace ngpb
