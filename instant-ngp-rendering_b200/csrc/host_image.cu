// The loader's other 8-bit image formats. The reference decodes every non-EXR dataset image with stb_image
// (load_stbi, src/common_host.cu:214-217; nerf_loader.cu:571-618), which picks the decoder from the file's content, not
// its name. JPEG goes to host_jpeg.cu (nvjpeg in the loader's default path) and PNG to host_png.cu; this file adds what is left of stb_image's list that datasets
// use - BMP, GIF (first frame), PSD (flattened composite), Softimage PIC, TGA and binary PNM - and the content sniffing in front of them. Output as everywhere: RGBA8, rows top to
// bottom, alpha 255 where the file has none. Host only.
#include "host_io.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>

namespace ngpb {

namespace {
template <typename T> T rd(const uint8_t* p) { T v; memcpy(&v, p, sizeof(T)); return v; }

[[noreturn]] void fail(const std::string& name, const std::string& why) {
	throw std::runtime_error{"Could not open image file '" + name + "': " + why};
}

// position of the lowest set bit and the number of set bits of a channel mask
int mask_shift(uint32_t m) { int s = 0; while (m && !(m & 1)) { m >>= 1; ++s; } return s; }
int mask_bits(uint32_t m) { int n = 0; while (m) { n += m & 1; m >>= 1; } return n; }
// a `bits`-wide channel value widened to 8 bits by bit replication (what stb_image does for 5- and 6-bit channels)
uint8_t widen(uint32_t v, int bits) {
	if (bits <= 0) return 0;
	if (bits >= 8) return (uint8_t)(v >> (bits - 8));
	uint32_t out = v << (8 - bits);
	for (int have = bits; have < 8; have += bits) out |= out >> have;
	return (uint8_t)out;
}

// ---- BMP: BITMAPCOREHEADER / INFOHEADER / V4 / V5; 1, 4, 8 bit palette, 16 / 24 / 32 bit, BI_RGB and BI_BITFIELDS ----
HostImage decode_bmp(const std::vector<uint8_t>& f, const std::string& name) {
	if (f.size() < 26) fail(name, "truncated BMP");
	const uint32_t data_at = rd<uint32_t>(&f[10]);
	const uint32_t hsz = rd<uint32_t>(&f[14]);
	if (hsz != 12 && hsz != 40 && hsz != 56 && hsz != 108 && hsz != 124) fail(name, "unknown BMP header");
	if (f.size() < 14 + (size_t)hsz) fail(name, "truncated BMP");
	int w, h, bpp;
	uint32_t compression = 0, mr = 0, mg = 0, mb = 0, ma = 0;
	if (hsz == 12) {
		w = rd<uint16_t>(&f[18]); h = rd<uint16_t>(&f[20]); bpp = rd<uint16_t>(&f[24]);
	} else {
		w = rd<int32_t>(&f[18]); h = rd<int32_t>(&f[22]); bpp = rd<uint16_t>(&f[28]);
		compression = rd<uint32_t>(&f[30]);
		if (compression == 1 || compression == 2) fail(name, "run-length coded BMPs are not supported");
		if (compression != 0 && compression != 3) fail(name, "unsupported BMP compression");
		if (compression == 3 || hsz >= 56) {
			// masks follow a 40-byte header, or are part of the larger ones
			if (f.size() < 14 + 40 + 12) fail(name, "truncated BMP");
			if (compression == 3) { mr = rd<uint32_t>(&f[54]); mg = rd<uint32_t>(&f[58]); mb = rd<uint32_t>(&f[62]); }
			if (hsz >= 56 && compression == 3) ma = rd<uint32_t>(&f[66]);
		}
	}
	const bool flip = h > 0;
	if (h < 0) h = -h;
	if (w <= 0 || h <= 0 || w > (1 << 16) || h > (1 << 16)) fail(name, "bad BMP size");
	if (bpp != 1 && bpp != 4 && bpp != 8 && bpp != 16 && bpp != 24 && bpp != 32) fail(name, "unsupported BMP bit depth");
	if (compression == 0) {
		if (bpp == 16) { mr = 31u << 10; mg = 31u << 5; mb = 31u; }
		else if (bpp == 32) { mr = 0xFFu << 16; mg = 0xFFu << 8; mb = 0xFFu; ma = 0xFFu << 24; }
	} else if (bpp != 16 && bpp != 32) {
		fail(name, "BMP bit fields need 16 or 32 bits per pixel");
	} else if (!mr || !mg || !mb) {
		fail(name, "bad BMP channel masks");
	}

	// palette: right after the header (and after the three masks of a 40-byte BI_BITFIELDS header)
	uint8_t palette[256][4] = {};
	if (bpp <= 8) {
		const size_t entry = hsz == 12 ? 3 : 4;
		size_t at = 14 + (size_t)hsz;
		size_t n = (data_at > at ? data_at - at : 0) / entry;
		if (n > 256) n = 256;
		if (at + n * entry > f.size()) fail(name, "truncated BMP palette");
		for (size_t i = 0; i < n; ++i, at += entry) {
			palette[i][2] = f[at]; palette[i][1] = f[at + 1]; palette[i][0] = f[at + 2]; palette[i][3] = 255;
		}
	}
	const size_t stride = (((size_t)w * bpp + 31) / 32) * 4;
	if (data_at > f.size() || (f.size() - data_at) / stride < (size_t)h) fail(name, "truncated BMP pixel data");

	HostImage img;
	img.width = w; img.height = h;
	img.rgba.resize((size_t)w * h * 4);
	const int rs = mask_shift(mr), gs = mask_shift(mg), bs = mask_shift(mb), as = mask_shift(ma);
	const int rb = mask_bits(mr), gb = mask_bits(mg), bb = mask_bits(mb), ab = mask_bits(ma);
	uint32_t any_alpha = 0;
	for (int y = 0; y < h; ++y) {
		const uint8_t* row = &f[data_at + stride * (size_t)y];
		uint8_t* out = &img.rgba[(size_t)(flip ? h - 1 - y : y) * w * 4];
		for (int x = 0; x < w; ++x, out += 4) {
			if (bpp <= 8) {
				const int per_byte = 8 / bpp;
				const int idx = (row[x / per_byte] >> (8 - bpp * (x % per_byte + 1))) & ((1 << bpp) - 1);
				memcpy(out, palette[idx], 4);
				out[3] = 255;
			} else if (bpp == 24) {
				out[0] = row[x * 3 + 2]; out[1] = row[x * 3 + 1]; out[2] = row[x * 3]; out[3] = 255;
			} else {
				const uint32_t v = bpp == 16 ? rd<uint16_t>(row + x * 2) : rd<uint32_t>(row + x * 4);
				out[0] = widen((v & mr) >> rs, rb);
				out[1] = widen((v & mg) >> gs, gb);
				out[2] = widen((v & mb) >> bs, bb);
				out[3] = ma ? widen((v & ma) >> as, ab) : 255;
				any_alpha |= out[3];
			}
		}
	}
	// a 32-bit file whose alpha bytes are all zero carries no alpha (stb_image's rule): opaque
	if (bpp == 32 && ma && any_alpha == 0) for (size_t i = 3; i < img.rgba.size(); i += 4) img.rgba[i] = 255;
	return img;
}

// ---- TGA: colour-mapped / true-colour / grey, plain or run-length coded, 8 / 15 / 16 / 24 / 32 bits ----
bool looks_like_tga(const std::vector<uint8_t>& f) {
	if (f.size() < 18) return false;
	const int cmap = f[1], type = f[2] & 7, bits = f[16];
	if (cmap > 1 || (f[2] & ~0x0B)) return false;
	if (type != 1 && type != 2 && type != 3) return false;
	if (cmap == 1) {
		if (type != 1) return false;
		const int eb = f[7];
		if (eb != 8 && eb != 15 && eb != 16 && eb != 24 && eb != 32) return false;
		if (bits != 8 && bits != 16) return false;
	} else {
		if (type == 1) return false;
		if (bits != 8 && bits != 15 && bits != 16 && bits != 24 && bits != 32) return false;
	}
	return rd<uint16_t>(&f[12]) > 0 && rd<uint16_t>(&f[14]) > 0;
}

// one stored TGA texel of `bits` bits -> RGBA8. 15 / 16-bit true colour is 1-5-5-5 with the top bit ignored and each 5-bit
// channel scaled as (v * 255) / 31, stb_image's rule (stbi__tga_read_rgb16) - not the bit replication BMP bit fields get
void tga_texel(const uint8_t* p, int bits, bool grey, uint8_t* out) {
	if (grey) {
		out[0] = out[1] = out[2] = p[0];
		out[3] = bits == 16 ? p[1] : 255;
	} else if (bits == 15 || bits == 16) {
		const uint32_t v = rd<uint16_t>(p);
		out[0] = (uint8_t)((((v >> 10) & 31) * 255) / 31); out[1] = (uint8_t)((((v >> 5) & 31) * 255) / 31); out[2] = (uint8_t)(((v & 31) * 255) / 31); out[3] = 255;
	} else {
		out[0] = p[2]; out[1] = p[1]; out[2] = p[0];
		out[3] = bits == 32 ? p[3] : 255;
	}
}

HostImage decode_tga(const std::vector<uint8_t>& f, const std::string& name) {
	if (!looks_like_tga(f)) fail(name, "not a supported TGA file");
	const int id_len = f[0], cmap = f[1], type = f[2] & 7;
	const bool rle = (f[2] & 8) != 0, grey = type == 3;
	const int cmap_first = rd<uint16_t>(&f[3]), cmap_len = rd<uint16_t>(&f[5]), cmap_bits = f[7];
	const int w = rd<uint16_t>(&f[12]), h = rd<uint16_t>(&f[14]), bits = f[16];
	const bool top_down = (f[17] >> 5) & 1;
	size_t at = 18 + (size_t)id_len;
	std::vector<uint8_t> palette;
	if (cmap) {
		// stb_image treats the colour map's "first entry index" as a count of BYTES to skip before the map and indexes the map
		// from 0 (stb_image.h stbi__tga_load: stbi__skip(s, tga_palette_start)); files in the wild have 0 there. Mirrored.
		at += (size_t)cmap_first;
		if (cmap_len == 0) fail(name, "bad palette");
		const size_t eb = (size_t)(cmap_bits + 7) / 8;
		if (at + eb * cmap_len > f.size()) fail(name, "truncated TGA palette");
		palette.resize((size_t)cmap_len * 4);
		for (int i = 0; i < cmap_len; ++i) tga_texel(&f[at + eb * i], cmap_bits, cmap_bits == 8, &palette[(size_t)i * 4]);
		at += eb * cmap_len;
	}
	const size_t texel = (size_t)(bits + 7) / 8;
	// a run packet expands at most 128 texels from 1 + texel bytes: refuse sizes the file cannot hold before allocating
	if (at > f.size() || (size_t)w * h > (f.size() - at) * (rle ? 128 : 1) / (rle ? 1 : texel)) fail(name, "truncated TGA pixel data");
	HostImage img;
	img.width = w; img.height = h;
	img.rgba.resize((size_t)w * h * 4);
	auto store = [&](size_t i, const uint8_t* p) {
		uint8_t* out = &img.rgba[i * 4];
		if (cmap) {
			int idx = bits == 8 ? p[0] : rd<uint16_t>(p);
			if (idx >= cmap_len) idx = 0;
			memcpy(out, &palette[(size_t)idx * 4], 4);
		} else {
			tga_texel(p, bits, grey, out);
		}
	};
	const size_t n = (size_t)w * h;
	if (!rle) {
		if (at + n * texel > f.size()) fail(name, "truncated TGA pixel data");
		for (size_t i = 0; i < n; ++i) store(i, &f[at + i * texel]);
	} else {
		size_t i = 0;
		while (i < n) {
			if (at >= f.size()) fail(name, "truncated TGA pixel data");
			const int packet = f[at++];
			const size_t count = (size_t)(packet & 127) + 1;
			if (i + count > n) fail(name, "corrupt TGA run");
			if (packet & 128) {
				if (at + texel > f.size()) fail(name, "truncated TGA pixel data");
				for (size_t k = 0; k < count; ++k) store(i++, &f[at]);
				at += texel;
			} else {
				if (at + texel * count > f.size()) fail(name, "truncated TGA pixel data");
				for (size_t k = 0; k < count; ++k, at += texel) store(i++, &f[at]);
			}
		}
	}
	if (!top_down) {
		std::vector<uint8_t> row((size_t)w * 4);
		for (int y = 0; y < h / 2; ++y) {
			uint8_t* a = &img.rgba[(size_t)y * w * 4];
			uint8_t* b = &img.rgba[(size_t)(h - 1 - y) * w * 4];
			memcpy(row.data(), a, row.size()); memcpy(a, b, row.size()); memcpy(b, row.data(), row.size());
		}
	}
	return img;
}

// ---- binary PNM: P5 (grey) and P6 (RGB), 8 or 16 bits per sample ----
// 16-bit samples are big-endian in the file; stb_image reads them as native words and keeps the upper half of those, which on
// the little-endian hosts the reference runs on is the file's SECOND byte - the low-order one. Mirrored, as the bar is the
// reference's texels (pinned by tests/test_host_io_cpu.py against the reference's stb_image).
HostImage decode_pnm(const std::vector<uint8_t>& f, const std::string& name) {
	size_t at = 2;
	auto number = [&]() {
		for (;;) {
			while (at < f.size() && (f[at] == ' ' || f[at] == '\t' || f[at] == '\n' || f[at] == '\r' || f[at] == '\f' || f[at] == '\v')) ++at;
			if (at < f.size() && f[at] == '#') { while (at < f.size() && f[at] != '\n' && f[at] != '\r') ++at; continue; }
			break;
		}
		if (at >= f.size() || f[at] < '0' || f[at] > '9') fail(name, "bad PNM header");
		long v = 0;
		while (at < f.size() && f[at] >= '0' && f[at] <= '9' && v < (1 << 24)) v = v * 10 + (f[at++] - '0');
		return (int)v;
	};
	const int channels = f[1] == '6' ? 3 : 1;
	const int w = number(), h = number(), maxval = number();
	++at;   // the single whitespace byte after maxval
	if (w <= 0 || h <= 0 || w > (1 << 16) || h > (1 << 16) || maxval <= 0 || maxval > 65535) fail(name, "bad PNM header");
	const size_t sample = maxval > 255 ? 2 : 1;
	const size_t n = (size_t)w * h;
	if (at > f.size() || (f.size() - at) / (sample * channels) < n) fail(name, "truncated PNM pixel data");
	HostImage img;
	img.width = w; img.height = h;
	img.rgba.resize(n * 4);
	for (size_t i = 0; i < n; ++i) {
		const uint8_t* p = &f[at + i * channels * sample];
		uint8_t* out = &img.rgba[i * 4];
		for (int c = 0; c < 3; ++c) out[c] = p[(channels == 3 ? c : 0) * sample + (sample - 1)];
		out[3] = 255;
	}
	return img;
}

// ---- GIF: the first frame, as stb_image's stbi_load returns it (stbi__gif_load_next, first_frame) ----
// Quirks kept because the bar is the reference's texels: pixels of the transparent index are left (0, 0, 0, 0); pixels outside
// the frame's rectangle take the background colour only when its index is > 0, alpha 255, and with red and blue swapped
// (stb_image copies its B-G-R-A palette entry as it is); a Graphic Control Extension whose block is not 4 bytes long is not
// skipped to its terminator. Bytes past the end of the file read as 0, which ends the raster early with what was drawn.
HostImage decode_gif(const std::vector<uint8_t>& f, const std::string& name) {
	size_t at = 0;
	auto get8 = [&]() -> int { return at < f.size() ? f[at++] : 0; };
	auto get16 = [&]() -> int { const int lo = get8(); return lo | (get8() << 8); };
	auto skip = [&](int n) { at = std::min(f.size(), at + (size_t)std::max(n, 0)); };
	if (f.size() < 13 || memcmp(f.data(), "GIF8", 4) != 0 || (f[4] != '7' && f[4] != '9') || f[5] != 'a') fail(name, "not GIF");
	at = 6;
	const int gw = get16(), gh = get16(), flags = get8(), bgindex = get8();
	get8();   // aspect ratio
	if (gw <= 0 || gh <= 0 || gw > (1 << 16) || gh > (1 << 16)) fail(name, "bad GIF size");
	uint8_t pal[256][4] = {}, lpal[256][4] = {};   // B, G, R, A as stb_image keeps them
	auto read_table = [&](uint8_t table[256][4], int n, int transparent) {
		for (int i = 0; i < n; ++i) {
			table[i][2] = (uint8_t)get8(); table[i][1] = (uint8_t)get8(); table[i][0] = (uint8_t)get8();
			table[i][3] = transparent == i ? 0 : 255;
		}
	};
	if (flags & 0x80) read_table(pal, 2 << (flags & 7), -1);
	if ((size_t)gw * gh > f.size() * 4096) fail(name, "GIF size implausible for the file");

	HostImage img;
	img.width = gw; img.height = gh;
	img.rgba.assign((size_t)gw * gh * 4, 0);
	std::vector<uint8_t> drawn((size_t)gw * gh, 0);
	int transparent = -1, eflags = 0;
	for (;;) {
		const int tag = get8();
		if (tag == 0x21) {
			const int ext = get8();
			int len;
			if (ext == 0xF9) {
				len = get8();
				if (len == 4) {
					eflags = get8();
					get16();   // delay
					if (transparent >= 0) pal[transparent][3] = 255;
					if (eflags & 0x01) { transparent = get8(); pal[transparent][3] = 0; }
					else { skip(1); transparent = -1; }
				} else {
					skip(len);
					continue;
				}
			}
			while ((len = get8()) != 0) skip(len);
		} else if (tag == 0x2C) {
			const int x = get16(), y = get16(), w = get16(), h = get16();
			if (x + w > gw || y + h > gh) fail(name, "bad Image Descriptor");
			const int lflags = get8();
			const size_t line = (size_t)gw * 4;
			const size_t start_x = (size_t)x * 4, start_y = (size_t)y * line, max_x = start_x + (size_t)w * 4, max_y = start_y + (size_t)h * line;
			size_t cur_x = start_x, cur_y = w == 0 ? max_y : start_y, step = (lflags & 0x40) ? 8 * line : line;
			int parse = (lflags & 0x40) ? 3 : 0;
			const uint8_t (*table)[4];
			if (lflags & 0x80) { read_table(lpal, 2 << (lflags & 7), (eflags & 0x01) ? transparent : -1); table = lpal; }
			else if (flags & 0x80) table = pal;
			else fail(name, "missing color table");

			auto emit = [&](uint8_t index) {
				if (cur_y >= max_y) return;
				const size_t idx = cur_x + cur_y;
				drawn[idx / 4] = 1;
				const uint8_t* c = table[index];
				if (c[3] > 128) { img.rgba[idx] = c[2]; img.rgba[idx + 1] = c[1]; img.rgba[idx + 2] = c[0]; img.rgba[idx + 3] = c[3]; }
				cur_x += 4;
				if (cur_x >= max_x) {
					cur_x = start_x;
					cur_y += step;
					while (cur_y >= max_y && parse > 0) {
						step = ((size_t)1 << parse) * line;
						cur_y = start_y + (step >> 1);
						--parse;
					}
				}
			};
			// LZW, code for code as stbi__process_gif_raster walks it
			struct Code { int16_t prefix; uint8_t first, suffix; };
			std::vector<Code> codes(8192);
			const int lzw_cs = get8();
			if (lzw_cs > 12) fail(name, "corrupt GIF raster");
			const int clear = 1 << lzw_cs;
			bool first = true;
			int codesize = lzw_cs + 1, codemask = (1 << codesize) - 1, avail = clear + 2, oldcode = -1, valid_bits = 0, len = 0;
			int32_t bits = 0;
			for (int c = 0; c < clear && c < 8192; ++c) codes[c] = {-1, (uint8_t)c, (uint8_t)c};
			std::vector<uint8_t> chain;
			bool done = false;
			while (!done) {
				if (valid_bits < codesize) {
					if (len == 0) {
						len = get8();
						if (len == 0) break;
					}
					--len;
					bits |= (int32_t)get8() << valid_bits;
					valid_bits += 8;
					continue;
				}
				const int code = bits & codemask;
				bits >>= codesize;
				valid_bits -= codesize;
				if (code == clear) {
					codesize = lzw_cs + 1; codemask = (1 << codesize) - 1; avail = clear + 2; oldcode = -1; first = false;
				} else if (code == clear + 1) {
					skip(len);
					while ((len = get8()) > 0) skip(len);
					done = true;
				} else if (code <= avail) {
					if (first) fail(name, "no clear code");
					if (oldcode >= 0) {
						if (avail + 1 > 8192) fail(name, "too many codes");
						Code& p = codes[avail++];
						p.prefix = (int16_t)oldcode;
						p.first = codes[oldcode].first;
						p.suffix = code == avail ? p.first : codes[code].first;
					} else if (code == avail) {
						fail(name, "illegal code in raster");
					}
					chain.clear();
					for (int c = code; c >= 0; c = codes[c].prefix) chain.push_back(codes[c].suffix);
					for (size_t k = chain.size(); k-- > 0;) emit(chain[k]);
					if ((avail & codemask) == 0 && avail <= 0x0FFF) { ++codesize; codemask = (1 << codesize) - 1; }
					oldcode = code;
				} else {
					fail(name, "illegal code in raster");
				}
			}
			if (bgindex > 0) {
				pal[bgindex][3] = 255;
				for (size_t p = 0; p < drawn.size(); ++p) if (!drawn[p]) memcpy(&img.rgba[p * 4], pal[bgindex], 4);
			}
			return img;
		} else if (tag == 0x3B) {
			fail(name, "GIF without an image");
		} else {
			fail(name, "unknown code");
		}
	}
}

// ---- PSD: the flattened composite at the end of the file, RGB mode, 8 or 16 bits, raw or PackBits, as stb_image reads it ----
// (stbi__psd_load): channels beyond the file's count are 0 (alpha 255); with four or more channels colours are taken off the white
// matte, pixel * (1 / a) + 255 * (1 - 1 / a) in float and cut to a byte the way the reference's x86 build cuts it; PackBits rows
// are read as bytes whatever the bit depth. Bytes past the end of the file read as 0.
HostImage decode_psd(const std::vector<uint8_t>& f, const std::string& name) {
	size_t at = 0;
	auto get8 = [&]() -> int { return at < f.size() ? f[at++] : 0; };
	auto get16 = [&]() -> int { const int hi = get8(); return (hi << 8) | get8(); };
	auto get32 = [&]() -> uint32_t { const uint32_t hi = (uint32_t)get16(); return (hi << 16) | (uint32_t)get16(); };
	auto skip = [&](uint32_t n) { at = (size_t)std::min<uint64_t>(f.size(), (uint64_t)at + n); };
	if (get32() != 0x38425053u) fail(name, "not PSD");
	if (get16() != 1) fail(name, "wrong version");
	skip(6);
	const int channel_count = get16();
	if (channel_count < 0 || channel_count > 16) fail(name, "wrong channel count");
	const uint32_t h = get32(), w = get32();
	if (w == 0 || h == 0 || w > (1u << 16) || h > (1u << 16)) fail(name, "bad PSD size");
	const int depth = get16();
	if (depth != 8 && depth != 16) fail(name, "unsupported bit depth");
	if (get16() != 3) fail(name, "wrong color format");
	skip(get32());   // colour mode data
	skip(get32());   // image resources
	skip(get32());   // layers and masks
	const int compression = get16();
	if (compression > 1) fail(name, "bad compression");
	const size_t n = (size_t)w * h;
	if (n > f.size() * 4096) fail(name, "PSD size implausible for the file");

	HostImage img;
	img.width = (int)w; img.height = (int)h;
	img.rgba.resize(n * 4);
	if (compression) skip((uint32_t)(h * (uint32_t)channel_count * 2));   // the per-row byte counts
	for (int c = 0; c < 4; ++c) {
		uint8_t* p = &img.rgba[c];
		if (c >= channel_count) {
			for (size_t i = 0; i < n; ++i, p += 4) *p = c == 3 ? 255 : 0;
		} else if (compression) {
			size_t count = 0;
			while (count < n) {
				const size_t left = n - count;
				int len = get8();
				if (len == 128) continue;
				if (len < 128) {
					++len;
					if ((size_t)len > left) fail(name, "bad RLE data");
					count += (size_t)len;
					for (; len; --len, p += 4) *p = (uint8_t)get8();
				} else {
					len = 257 - len;
					if ((size_t)len > left) fail(name, "bad RLE data");
					const uint8_t v = (uint8_t)get8();
					count += (size_t)len;
					for (; len; --len, p += 4) *p = v;
				}
				if (at >= f.size() && count < n) { for (; count < n; ++count, p += 4) *p = 0; }   // zeros past the end decode as literals of 0
			}
		} else if (depth == 16) {
			for (size_t i = 0; i < n; ++i, p += 4) *p = (uint8_t)(get16() >> 8);
		} else {
			for (size_t i = 0; i < n; ++i, p += 4) *p = (uint8_t)get8();
		}
	}
	if (channel_count >= 4) {
		for (size_t i = 0; i < n; ++i) {
			uint8_t* px = &img.rgba[i * 4];
			if (px[3] != 0 && px[3] != 255) {
				const float a = px[3] / 255.0f, ra = 1.0f / a, inv_a = 255.0f * (1 - ra);
				for (int c = 0; c < 3; ++c) px[c] = (uint8_t)(int)(px[c] * ra + inv_a);
			}
		}
	}
	return img;
}

// ---- Softimage PIC, as stb_image reads it (stbi__pic_load): up to 10 channel packets of 8 bits, raw / pure RLE / mixed RLE ----
bool looks_like_pic(const std::vector<uint8_t>& f) {
	return f.size() >= 92 && f[0] == 0x53 && f[1] == 0x80 && f[2] == 0xF6 && f[3] == 0x34 && memcmp(&f[88], "PICT", 4) == 0;
}

HostImage decode_pic(const std::vector<uint8_t>& f, const std::string& name) {
	size_t at = 92;
	auto eof = [&]() { return at >= f.size(); };
	auto get8 = [&]() -> int { return at < f.size() ? f[at++] : 0; };
	auto get16 = [&]() -> int { const int hi = get8(); return (hi << 8) | get8(); };
	const int w = get16(), h = get16();
	if (eof()) fail(name, "file too short (pic header)");
	if (w <= 0 || h <= 0) fail(name, "bad PIC size");
	at += 8;   // ratio, fields, pad
	struct Packet { int type, channel; };
	std::vector<Packet> packets;
	int chained;
	do {
		if (packets.size() == 10) fail(name, "too many packets");
		chained = get8();
		const int size = get8(), type = get8(), channel = get8();
		if (eof()) fail(name, "file too short (reading packets)");
		if (size != 8) fail(name, "packet isn't 8bpp");
		packets.push_back({type, channel});
	} while (chained);
	if ((size_t)w * h > f.size() * 65536) fail(name, "PIC size implausible for the file");

	HostImage img;
	img.width = w; img.height = h;
	img.rgba.assign((size_t)w * h * 4, 0xFF);
	auto readval = [&](int channel, uint8_t* dest) {
		for (int i = 0, mask = 0x80; i < 4; ++i, mask >>= 1) {
			if (channel & mask) {
				if (eof()) fail(name, "PIC file too short");
				dest[i] = (uint8_t)get8();
			}
		}
	};
	auto copyval = [](int channel, uint8_t* dest, const uint8_t* src) {
		for (int i = 0, mask = 0x80; i < 4; ++i, mask >>= 1) if (channel & mask) dest[i] = src[i];
	};
	for (int y = 0; y < h; ++y) {
		for (const Packet& p : packets) {
			uint8_t* dest = &img.rgba[(size_t)y * w * 4];
			if (p.type == 0) {
				for (int x = 0; x < w; ++x, dest += 4) readval(p.channel, dest);
			} else if (p.type == 1) {
				for (int left = w; left > 0;) {
					int count = get8();
					if (eof()) fail(name, "file too short (pure read count)");
					if (count > left) count = (uint8_t)left;
					uint8_t value[4] = {};
					readval(p.channel, value);
					for (int i = 0; i < count; ++i, dest += 4) copyval(p.channel, dest, value);
					left -= count;
				}
			} else if (p.type == 2) {
				for (int left = w; left > 0;) {
					int count = get8();
					if (eof()) fail(name, "file too short (mixed read count)");
					if (count >= 128) {
						count = count == 128 ? get16() : count - 127;
						if (count > left) fail(name, "scanline overrun");
						uint8_t value[4] = {};
						readval(p.channel, value);
						for (int i = 0; i < count; ++i, dest += 4) copyval(p.channel, dest, value);
					} else {
						++count;
						if (count > left) fail(name, "scanline overrun");
						for (int i = 0; i < count; ++i, dest += 4) readval(p.channel, dest);
					}
					left -= count;
				}
			} else {
				fail(name, "packet has bad compression type");
			}
		}
	}
	return img;
}

// ---- Radiance HDR (RGBE), as stb_image reads it: stbi_loadf gives the floats (the image mode's path, load_stbi_gpu,
// src/common_host.cu:268-299), stbi_load turns them into bytes with gamma 1 / 2.2 (what a dataset frame would get) ----
bool looks_like_hdr(const std::vector<uint8_t>& f) {
	return (f.size() >= 11 && memcmp(f.data(), "#?RADIANCE\n", 11) == 0) || (f.size() >= 7 && memcmp(f.data(), "#?RGBE\n", 7) == 0);
}

void rgbe_to_float(float* out, const uint8_t* in) {
	if (in[3] != 0) {
		const float f1 = (float)ldexp(1.0f, in[3] - (int)(128 + 8));
		out[0] = in[0] * f1; out[1] = in[1] * f1; out[2] = in[2] * f1;
	} else {
		out[0] = out[1] = out[2] = 0;
	}
	out[3] = 1;
}
} // namespace

HostImageF decode_hdr(const std::vector<uint8_t>& f, const std::string& name) {
	size_t at = 0;
	auto eof = [&]() { return at >= f.size(); };
	auto get8 = [&]() -> int { return at < f.size() ? f[at++] : 0; };
	auto token = [&]() {   // one header line, at most 1023 characters of it
		std::string t;
		char c = (char)get8();
		while (!eof() && c != '\n') {
			t += c;
			if (t.size() == 1023) { while (!eof() && get8() != '\n') {} break; }
			c = (char)get8();
		}
		return t;
	};
	const std::string magic = token();
	if (magic != "#?RADIANCE" && magic != "#?RGBE") fail(name, "not HDR");
	bool valid = false;
	for (;;) {
		const std::string t = token();
		if (t.empty()) break;
		if (t == "FORMAT=32-bit_rle_rgbe") valid = true;
	}
	if (!valid) fail(name, "unsupported format");
	const std::string dims = token();
	if (dims.compare(0, 3, "-Y ") != 0) fail(name, "unsupported data layout");
	char* rest = nullptr;
	const int height = (int)strtol(dims.c_str() + 3, &rest, 10);
	while (*rest == ' ') ++rest;
	if (strncmp(rest, "+X ", 3) != 0) fail(name, "unsupported data layout");
	const int width = (int)strtol(rest + 3, nullptr, 10);
	if (width <= 0 || height <= 0 || width > (1 << 24) || height > (1 << 24)) fail(name, "bad HDR size");
	if ((size_t)width * height > f.size() * 4096) fail(name, "HDR size implausible for the file");

	HostImageF img;
	img.width = width; img.height = height;
	img.rgba.assign((size_t)width * height * 4, 0.0f);
	const size_t n = (size_t)width * height;
	auto flat_from = [&](size_t first) {
		for (size_t i = first; i < n; ++i) {
			uint8_t rgbe[4];
			for (int k = 0; k < 4; ++k) rgbe[k] = (uint8_t)get8();
			rgbe_to_float(&img.rgba[i * 4], rgbe);
		}
	};
	if (width < 8 || width >= 32768) {
		flat_from(0);
		return img;
	}
	std::vector<uint8_t> scanline((size_t)width * 4);
	for (int j = 0; j < height; ++j) {
		const int c1 = get8(), c2 = get8();
		int len = get8();
		if (c1 != 2 || c2 != 2 || (len & 0x80)) {
			// not a run-length coded line: stb_image takes these four bytes as the FIRST pixel of the picture and reads all the
			// others flat from here, whichever line it was on
			const uint8_t rgbe[4] = {(uint8_t)c1, (uint8_t)c2, (uint8_t)len, (uint8_t)get8()};
			rgbe_to_float(&img.rgba[0], rgbe);
			flat_from(1);
			return img;
		}
		len = (len << 8) | get8();
		if (len != width) fail(name, "invalid decoded scanline length");
		for (int k = 0; k < 4; ++k) {
			for (int i = 0; i < width;) {
				const int left = width - i;
				int count = get8();
				if (count > 128) {
					const uint8_t value = (uint8_t)get8();
					count -= 128;
					if (count == 0 || count > left) fail(name, "bad RLE data in HDR");
					for (int z = 0; z < count; ++z) scanline[(size_t)i++ * 4 + k] = value;
				} else {
					if (count == 0 || count > left) fail(name, "bad RLE data in HDR");
					for (int z = 0; z < count; ++z) scanline[(size_t)i++ * 4 + k] = (uint8_t)get8();
				}
			}
		}
		for (int i = 0; i < width; ++i) rgbe_to_float(&img.rgba[((size_t)j * width + i) * 4], &scanline[(size_t)i * 4]);
	}
	return img;
}

namespace {
// stbi__hdr_to_ldr: pow(v, 1 / 2.2) * 255 + 0.5, cut to [0, 255]; alpha linear
HostImage hdr_to_bytes(const HostImageF& src) {
	HostImage img;
	img.width = src.width; img.height = src.height;
	img.rgba.resize(src.rgba.size());
	for (size_t i = 0; i < src.rgba.size(); ++i) {
		float z = (i & 3) == 3 ? src.rgba[i] * 255 + 0.5f : (float)pow((double)(src.rgba[i] * 1.0f), (double)(1.0f / 2.2f)) * 255 + 0.5f;
		if (z < 0) z = 0;
		if (z > 255) z = 255;
		img.rgba[i] = (uint8_t)(int)z;
	}
	return img;
}
} // namespace

bool is_hdr_file(const std::vector<uint8_t>& f) { return looks_like_hdr(f); }
bool is_jpeg_file(const std::vector<uint8_t>& f) { return f.size() > 3 && f[0] == 0xFF && f[1] == 0xD8 && f[2] == 0xFF; }

HostImage decode_image_rgba8(const std::vector<uint8_t>& f, const std::string& name) {
	static const uint8_t png_magic[8] = {0x89, 'P', 'N', 'G', '\r', '\n', 0x1A, '\n'};
	if (f.size() >= 8 && memcmp(f.data(), png_magic, 8) == 0) return decode_png(f, name);
	if (is_jpeg_file(f)) return decode_jpeg(f, name);
	if (f.size() >= 4 && memcmp(f.data(), "GIF8", 4) == 0) return decode_gif(f, name);
	if (f.size() >= 2 && f[0] == 'B' && f[1] == 'M') return decode_bmp(f, name);
	if (f.size() >= 4 && memcmp(f.data(), "8BPS", 4) == 0) return decode_psd(f, name);
	if (looks_like_pic(f)) return decode_pic(f, name);
	if (looks_like_hdr(f)) return hdr_to_bytes(decode_hdr(f, name));
	if (f.size() >= 3 && f[0] == 'P' && (f[1] == '5' || f[1] == '6')) return decode_pnm(f, name);
	if (looks_like_tga(f)) return decode_tga(f, name);
	fail(name, "unknown image type (JPEG, PNG, BMP, GIF, PSD, PIC, TGA, binary PNM, Radiance HDR and EXR are read)");
}

} // namespace ngpb
