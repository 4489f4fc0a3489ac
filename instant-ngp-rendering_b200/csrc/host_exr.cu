// OpenEXR reader for the loader (the reference reads EXR through tinyexr, src/tinyexr_wrapper.cu, which is not part of
// this image): single-part scanline and tiled (full-resolution level) files, HALF and FLOAT channels, compression NONE / RLE / ZIPS / ZIP / PIZ - the set
// tinyexr decodes (data/image/albert.exr is FLOAT RGBA, ZIP; PIZ is what most DCC tools write by default). PXR24 / B44 /
// DWA, deep and multi-part files are refused with a message. Output: float RGBA, rows top to bottom; which channels land
// where follows the reference's two readers (ExrChannelRule in host_io.h). Host only; zlib does the inflate; the PIZ stages (Huffman, 2-D Haar
// wavelet, value LUT) follow the OpenEXR file-format description.
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

// ZIP / RLE post-pass: undo the byte-delta predictor, then interleave the two halves back into byte pairs.
void unpredict_and_interleave(std::vector<uint8_t>& tmp, uint8_t* raw) {
	const size_t n = tmp.size();
	for (size_t i = 1; i < n; ++i) tmp[i] = (uint8_t)(tmp[i - 1] + tmp[i] - 128);
	const size_t half = (n + 1) / 2;
	for (size_t i = 0; i < n; ++i) raw[i] = (i & 1) ? tmp[half + i / 2] : tmp[i / 2];
}

// RLE chunk: a signed count byte, negative = that many literal bytes, otherwise the next byte count + 1 times.
bool rle_expand(const uint8_t* src, size_t n_src, std::vector<uint8_t>& out) {
	size_t o = 0, i = 0;
	while (i < n_src) {
		const int count = (int8_t)src[i++];
		if (count < 0) {
			const size_t n = (size_t)-count;
			if (i + n > n_src || o + n > out.size()) return false;
			memcpy(&out[o], &src[i], n);
			i += n; o += n;
		} else {
			const size_t n = (size_t)count + 1;
			if (i >= n_src || o + n > out.size()) return false;
			memset(&out[o], src[i++], n);
			o += n;
		}
	}
	return o == out.size();
}

// ---- PIZ ----
// Canonical Huffman code over the 16-bit alphabet plus one run-length symbol. A code table entry is (code << 6) | length.
constexpr int kHufSymbols = (1 << 16) + 1;
constexpr int kHufDecBits = 14;

struct HufReader {
	const uint8_t* p;
	const uint8_t* end;
	uint64_t bits = 0;
	int n_bits = 0;
	bool refill() {
		if (p >= end) return false;
		bits = (bits << 8) | *p++;
		n_bits += 8;
		return true;
	}
	// the next n bits (n <= 16), or -1 at the end of the input
	int take(int n) {
		while (n_bits < n) if (!refill()) return -1;
		n_bits -= n;
		return (int)((bits >> n_bits) & ((1u << n) - 1));
	}
};

// Code lengths of symbols [first, last], 6 bits each; 59..62 are short runs of zero lengths, 63 a long run (8 more bits).
bool huf_read_lengths(HufReader& r, int first, int last, std::vector<uint64_t>& table) {
	for (int s = first; s <= last; ++s) {
		const int l = r.take(6);
		if (l < 0) return false;
		int zeros = 0;
		if (l == 63) {
			const int more = r.take(8);
			if (more < 0) return false;
			zeros = more + 6;
		} else if (l >= 59) {
			zeros = l - 59 + 2;
		} else {
			table[s] = (uint64_t)l;
			continue;
		}
		if (s + zeros > last + 1) return false;
		s += zeros - 1;
	}
	return true;
}

// Lengths -> canonical codes: within a length codes count upwards in symbol order, shorter lengths take the numerically larger prefixes.
void huf_assign_codes(std::vector<uint64_t>& table) {
	uint64_t count[59] = {};
	for (uint64_t e : table) ++count[e];
	uint64_t code = 0;
	for (int l = 58; l > 0; --l) {
		const uint64_t next = (code + count[l]) >> 1;
		count[l] = code;
		code = next;
	}
	for (uint64_t& e : table) {
		const uint64_t l = e;
		if (l > 0) e = l | (count[l]++ << 6);
	}
}

// Decodes n_out symbols. Codes of up to 14 bits resolve in one lookup of the next 14 bits; longer ones are searched among
// the symbols sharing that 14-bit prefix. The run-length symbol `run_symbol` is followed by an 8-bit count and repeats the
// previous output value.
bool huf_decode(const std::vector<uint64_t>& table, int first, int last, const uint8_t* src, size_t n_src, uint64_t n_bits_total,
		int run_symbol, uint16_t* out, size_t n_out) {
	std::vector<uint32_t> primary(1u << kHufDecBits, 0);   // length | symbol << 8, 0 = a long code starts here
	std::map<uint32_t, std::vector<int>> longer;
	for (int s = first; s <= last; ++s) {
		const uint64_t code = table[s] >> 6;
		const int l = (int)(table[s] & 63);
		if (l == 0) continue;
		if (code >> l) return false;
		if (l > kHufDecBits) {
			const uint32_t prefix = (uint32_t)(code >> (l - kHufDecBits));
			if (primary[prefix]) return false;
			longer[prefix].push_back(s);
		} else {
			const uint32_t base = (uint32_t)(code << (kHufDecBits - l));
			for (uint32_t i = 0; i < (1u << (kHufDecBits - l)); ++i) {
				if (primary[base + i] || longer.count(base + i)) return false;
				primary[base + i] = (uint32_t)l | ((uint32_t)s << 8);
			}
		}
	}
	if ((n_bits_total + 7) / 8 > n_src) return false;
	HufReader r{src, src + (n_bits_total + 7) / 8};
	size_t o = 0;
	auto emit = [&](int symbol) {
		if (symbol == run_symbol) {
			if (r.n_bits < 8 && !r.refill()) return false;
			r.n_bits -= 8;
			const size_t run = (size_t)((r.bits >> r.n_bits) & 0xFF);
			if (o == 0 || o + run > n_out) return false;
			const uint16_t v = out[o - 1];
			for (size_t i = 0; i < run; ++i) out[o++] = v;
			return true;
		}
		if (o >= n_out) return false;
		out[o++] = (uint16_t)symbol;
		return true;
	};
	while (r.p < r.end) {
		r.refill();
		while (r.n_bits >= kHufDecBits) {
			const uint32_t window = (uint32_t)(r.bits >> (r.n_bits - kHufDecBits)) & ((1u << kHufDecBits) - 1);
			const uint32_t e = primary[window];
			if (e) {
				r.n_bits -= (int)(e & 0xFF);
				if (!emit((int)(e >> 8))) return false;
				continue;
			}
			const auto bucket = longer.find(window);
			if (bucket == longer.end()) return false;
			bool found = false;
			for (int s : bucket->second) {
				const int l = (int)(table[s] & 63);
				while (r.n_bits < l && r.refill()) {}
				if (r.n_bits >= l && (table[s] >> 6) == ((r.bits >> (r.n_bits - l)) & ((1ull << l) - 1))) {
					r.n_bits -= l;
					if (!emit(s)) return false;
					found = true;
					break;
				}
			}
			if (!found) return false;
		}
	}
	// the last byte carries padding bits; what is left are short codes
	const int padding = (int)((8 - n_bits_total) & 7);
	r.bits >>= padding;
	r.n_bits -= padding;
	while (r.n_bits > 0) {
		const uint32_t e = primary[(uint32_t)(r.bits << (kHufDecBits - r.n_bits)) & ((1u << kHufDecBits) - 1)];
		if (!e) return false;
		r.n_bits -= (int)(e & 0xFF);
		if (r.n_bits < 0 || !emit((int)(e >> 8))) return false;
	}
	return o == n_out;
}

bool huf_uncompress(const uint8_t* src, size_t n_src, uint16_t* out, size_t n_out) {
	if (n_src == 0) return n_out == 0;
	if (n_src < 20) return false;
	const uint32_t first = rd<uint32_t>(src), last = rd<uint32_t>(src + 4), n_bits = rd<uint32_t>(src + 12);
	if (first >= (uint32_t)kHufSymbols || last >= (uint32_t)kHufSymbols || first > last) return false;
	std::vector<uint64_t> table(kHufSymbols, 0);
	HufReader r{src + 20, src + n_src};
	if (!huf_read_lengths(r, (int)first, (int)last, table)) return false;
	huf_assign_codes(table);
	return huf_decode(table, (int)first, (int)last, r.p, (size_t)(r.end - r.p), n_bits, (int)last, out, n_out);
}

// One inverse Haar step on a (low, high) pair. Below 2^14 distinct values the transform is exact in 16 bits signed;
// otherwise it works modulo 2^16 around an offset of 2^15.
inline void wavelet_pair(bool small_range, uint16_t low, uint16_t high, uint16_t& a, uint16_t& b) {
	if (small_range) {
		const int h = (int16_t)high;
		const int av = (int16_t)low + (h & 1) + (h >> 1);
		a = (uint16_t)(int16_t)av;
		b = (uint16_t)(int16_t)(av - h);
	} else {
		const int bv = ((int)low - ((int)high >> 1)) & 0xFFFF;
		a = (uint16_t)(((int)high + bv - (1 << 15)) & 0xFFFF);
		b = (uint16_t)bv;
	}
}

// Inverse 2-D wavelet over an nx * ny plane of 16-bit words with element strides ox / oy, coarsest level first.
void wavelet_decode(uint16_t* plane, int nx, int ox, int ny, int oy, uint16_t max_value) {
	const bool small_range = max_value < (1 << 14);
	const int n = std::min(nx, ny);
	int p = 1;
	while (p <= n) p <<= 1;
	p >>= 1;
	int p2 = p;
	p >>= 1;
	for (; p >= 1; p2 = p, p >>= 1) {
		const ptrdiff_t oy1 = (ptrdiff_t)oy * p, ox1 = (ptrdiff_t)ox * p;
		int y = 0;
		for (; y + p2 <= ny; y += p2) {
			uint16_t* row = plane + (ptrdiff_t)oy * y;
			int x = 0;
			for (; x + p2 <= nx; x += p2) {
				uint16_t* px = row + (ptrdiff_t)ox * x;
				uint16_t i00, i01, i10, i11;
				wavelet_pair(small_range, px[0], px[oy1], i00, i10);
				wavelet_pair(small_range, px[ox1], px[oy1 + ox1], i01, i11);
				wavelet_pair(small_range, i00, i01, px[0], px[ox1]);
				wavelet_pair(small_range, i10, i11, px[oy1], px[oy1 + ox1]);
			}
			if (nx & p) {   // odd column: vertical pair only
				uint16_t* px = row + (ptrdiff_t)ox * x;
				uint16_t a, b;
				wavelet_pair(small_range, px[0], px[oy1], a, b);
				px[0] = a; px[oy1] = b;
			}
		}
		if (ny & p) {       // odd row: horizontal pairs only
			uint16_t* row = plane + (ptrdiff_t)oy * y;
			for (int x = 0; x + p2 <= nx; x += p2) {
				uint16_t* px = row + (ptrdiff_t)ox * x;
				uint16_t a, b;
				wavelet_pair(small_range, px[0], px[ox1], a, b);
				px[0] = a; px[ox1] = b;
			}
		}
	}
}

// A PIZ chunk: [min, max] of the used-value bitmap, the bitmap bytes in that range, the Huffman stream length, the stream.
// Decoded words are wavelet coefficients per channel plane (FLOAT channels are two interleaved 16-bit planes); after the
// inverse transform the value LUT maps dense ranks back to the 16-bit values that occur, and planes are re-cut into lines.
bool piz_uncompress(const uint8_t* src, size_t n_src, const std::vector<ExrChannel>& channels, int w, int lines, uint8_t* raw, size_t n_raw) {
	if (n_src < 4) return false;
	std::vector<uint8_t> bitmap(8192, 0);
	const uint16_t lo = rd<uint16_t>(src), hi = rd<uint16_t>(src + 2);
	size_t pos = 4;
	if (hi >= 8192) return false;
	if (lo <= hi) {
		const size_t n = (size_t)hi - lo + 1;
		if (pos + n > n_src) return false;
		memcpy(&bitmap[lo], src + pos, n);
		pos += n;
	}
	std::vector<uint16_t> lut(1 << 16, 0);
	int n_values = 0;
	for (int v = 0; v < (1 << 16); ++v) if (v == 0 || (bitmap[v >> 3] & (1 << (v & 7)))) lut[n_values++] = (uint16_t)v;
	const uint16_t max_value = (uint16_t)(n_values - 1);

	if (pos + 4 > n_src) return false;
	const int32_t n_huf = rd<int32_t>(src + pos);
	pos += 4;
	if (n_huf < 0 || pos + (size_t)n_huf > n_src) return false;
	const size_t n_words = n_raw / 2;
	std::vector<uint16_t> words(n_words);
	if (!huf_uncompress(src + pos, (size_t)n_huf, words.data(), n_words)) return false;

	std::vector<size_t> plane_start(channels.size());
	size_t at = 0;
	for (size_t c = 0; c < channels.size(); ++c) {
		const int words_per_pixel = channels[c].pixel_type == 1 ? 1 : 2;
		plane_start[c] = at;
		for (int j = 0; j < words_per_pixel; ++j) wavelet_decode(&words[at + j], w, words_per_pixel, lines, w * words_per_pixel, max_value);
		at += (size_t)w * lines * words_per_pixel;
	}
	if (at != n_words) return false;
	for (uint16_t& v : words) v = lut[v];
	uint16_t* out = (uint16_t*)raw;
	for (int l = 0; l < lines; ++l) {
		for (size_t c = 0; c < channels.size(); ++c) {
			const size_t n = (size_t)w * (channels[c].pixel_type == 1 ? 1 : 2);
			memcpy(out, &words[plane_start[c] + (size_t)l * n], n * 2);
			out += n;
		}
	}
	return true;
}
} // namespace

HostImageF decode_exr(const std::vector<uint8_t>& file, const std::string& name, ExrChannelRule rule) {
	auto fail = [&](const std::string& why) -> void { throw std::runtime_error{"Could not load image file '" + name + "': " + why}; };
	if (file.size() < 8 || rd<uint32_t>(file.data()) != 20000630u) fail("not an OpenEXR file");
	const uint32_t version = rd<uint32_t>(file.data() + 4);
	if ((version & 0xFF) != 2) fail("unsupported OpenEXR version");
	const bool tiled = (version & 0x200) != 0;
	if (version & 0x1800) fail("deep / multi-part files are not supported");

	std::vector<ExrChannel> channels;
	int compression = -1, xmin = 0, ymin = 0, xmax = -1, ymax = -1;
	uint32_t tile_w = 0, tile_h = 0, tile_mode = 0;
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
		else if (attr == "tiles" && size >= 9) { tile_w = rd<uint32_t>(v); tile_h = rd<uint32_t>(v + 4); tile_mode = v[8]; }
		else if (attr == "dataWindow") { xmin = rd<int32_t>(v); ymin = rd<int32_t>(v + 4); xmax = rd<int32_t>(v + 8); ymax = rd<int32_t>(v + 12); }
		pos += size;
	}
	const int w = xmax - xmin + 1, h = ymax - ymin + 1;
	if (channels.empty() || w <= 0 || h <= 0 || w > (1 << 16) || h > (1 << 16)) fail("missing channels or data window");
	int lines_per_block = 0;
	switch (compression) {
		case 0: case 1: case 2: lines_per_block = 1; break;
		case 3: lines_per_block = 16; break;
		case 4: lines_per_block = 32; break;
		default: fail("compression type " + std::to_string(compression) + " is not supported (NONE, RLE, ZIPS, ZIP and PIZ are)");
	}
	size_t bytes_per_pixel_column = 0;   // bytes one pixel takes over all channels
	for (const ExrChannel& ch : channels) {
		if (ch.pixel_type != 1 && ch.pixel_type != 2) fail("UINT channels are not supported");
		bytes_per_pixel_column += ch.pixel_type == 1 ? 2 : 4;
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
	if (rule == ExrChannelRule::Image) {
		// LoadEXR (tinyexr.h): a one-channel file of any name fills all four outputs with it; otherwise R, G and B must exist
		if (channels.size() == 1) {
			grey = 0;
			target[0] = -1;
			has_alpha = true;
		} else {
			for (int k = 0; k < 3; ++k) {
				bool found = false;
				for (size_t c = 0; c < channels.size(); ++c) found = found || target[c] == k;
				if (!found) fail(std::string{"RGB"[k]} + " channel not found");
			}
		}
	} else {
		// load_exr_to_gpu (src/tinyexr_wrapper.cu:192-244): channels by name, absent ones stay 0, one pixel type for all
		for (const ExrChannel& ch : channels) if (ch.pixel_type != channels[0].pixel_type) fail("Can't handle EXR images with mixed channel types");
	}
	(void)has_rgb;

	// tiled files (tinyexr reads the full-resolution level of them): one chunk per tile, the tiles of lower levels after them
	if (tiled && (tile_w == 0 || tile_h == 0 || tile_w > (1u << 16) || tile_h > (1u << 16))) fail("bad tile description");
	int n_blocks = (h + lines_per_block - 1) / lines_per_block;
	if (tiled) {
		const uint64_t nx = ((uint64_t)w + tile_w - 1) / tile_w, ny = ((uint64_t)h + tile_h - 1) / tile_h;
		(void)tile_mode;   // mip / rip files: tinyexr reads the first nx * ny offsets as the full-resolution level, and so does this
		n_blocks = (int)(nx * ny);
	}
	if (pos + (size_t)n_blocks * 8 > file.size()) fail("truncated offset table");
	HostImageF img;
	img.width = w; img.height = h;
	img.rgba.assign((size_t)w * h * 4, 0.0f);
	if (!has_alpha) for (size_t i = 0; i < (size_t)w * h; ++i) img.rgba[i * 4 + 3] = 1.0f;
	std::vector<uint8_t> raw, tmp;
	// one chunk = one rectangle of the image: a block of scanlines, or a tile (clipped at the right and bottom edges)
	for (int b = 0; b < n_blocks; ++b) {
		const uint64_t off = rd<uint64_t>(&file[pos + (size_t)b * 8]);
		const size_t head = tiled ? 20 : 8;
		if (off + head > file.size()) fail("bad chunk offset");
		int x0 = 0, y0, bw = w, lines;
		if (tiled) {
			const int tx = rd<int32_t>(&file[off]), ty = rd<int32_t>(&file[off + 4]);
			if (rd<int32_t>(&file[off + 8]) != 0 || rd<int32_t>(&file[off + 12]) != 0) fail("tiles of a lower-resolution level are not supported");
			if (tx < 0 || ty < 0 || (uint64_t)tx * tile_w >= (uint64_t)w || (uint64_t)ty * tile_h >= (uint64_t)h) fail("bad tile");
			x0 = tx * (int)tile_w; y0 = ty * (int)tile_h;
			bw = std::min((int)tile_w, w - x0);
			lines = std::min((int)tile_h, h - y0);
		} else {
			y0 = rd<int32_t>(&file[off]) - ymin;
			if (y0 < 0 || y0 >= h) fail("bad chunk");
			lines = std::min(lines_per_block, h - y0);
		}
		const uint32_t packed = rd<uint32_t>(&file[off + head - 4]);
		if (off + head + packed > file.size()) fail("bad chunk");
		const size_t unpacked = bytes_per_pixel_column * (size_t)bw * lines;
		const uint8_t* src = &file[off + head];
		raw.resize(unpacked);
		if (compression == 0 || packed == unpacked) {
			if (packed != unpacked) fail("bad uncompressed chunk size");
			memcpy(raw.data(), src, unpacked);
		} else if (compression == 4) {
			if (!piz_uncompress(src, packed, channels, bw, lines, raw.data(), unpacked)) fail("corrupt PIZ chunk");
		} else if (compression == 1) {
			tmp.resize(unpacked);
			if (!rle_expand(src, packed, tmp)) fail("corrupt RLE chunk");
			unpredict_and_interleave(tmp, raw.data());
		} else {
			tmp.resize(unpacked);
			uLongf got = (uLongf)unpacked;
			if (uncompress(tmp.data(), &got, src, packed) != Z_OK || got != unpacked) fail("corrupt ZIP chunk");
			unpredict_and_interleave(tmp, raw.data());
		}
		const uint8_t* p = raw.data();
		for (int l = 0; l < lines; ++l) {
			float* out = &img.rgba[((size_t)(y0 + l) * w + x0) * 4];
			for (size_t c = 0; c < channels.size(); ++c) {
				const bool is_half = channels[c].pixel_type == 1;
				const int t = (int)c == grey ? 4 : target[c];
				if (t >= 0) {
					for (int x = 0; x < bw; ++x) {
						const float v = is_half ? __half2float(rd<__half>(p + (size_t)x * 2)) : rd<float>(p + (size_t)x * 4);
						if (t == 4) { out[x * 4] = out[x * 4 + 1] = out[x * 4 + 2] = out[x * 4 + 3] = v; }
						else out[x * 4 + t] = v;
					}
				}
				p += (size_t)bw * (is_half ? 2 : 4);
			}
		}
	}
	return img;
}

} // namespace ngpb
