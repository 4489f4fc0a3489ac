// PNG decoder on zlib for the dataset loader (the reference decodes with stb_image, src/nerf_loader.cu:541-560, which
// is not linked here; the reference's vendored copy, compiled for the host, pins this file in tests/test_host_io_cpu.py):
// plain and Adam7-interlaced files, 8- and 16-bit samples (16 -> 8 by the high byte, as stb_image's conversion), grey /
// grey-alpha / RGB / RGBA / palette with tRNS, bit depths 1 / 2 / 4 for grey and palette. Output is 8-bit RGBA, rows top
// to bottom. Host only.
#include "host_io.h"

#include <zlib.h>

#include <cstring>
#include <stdexcept>

namespace ngpb {

static uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

static uint8_t paeth(int a, int b, int c) {
	const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
	return (uint8_t)((pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c));
}

HostImage decode_png(const std::vector<uint8_t>& file, const std::string& name) {
	static const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
	auto fail = [&](const std::string& why) -> void { throw std::runtime_error{"Could not decode PNG '" + name + "': " + why}; };
	if (file.size() < 8 || memcmp(file.data(), sig, 8) != 0) fail("not a PNG file");

	uint32_t w = 0, h = 0;
	int depth = 0, color = 0, interlace = 0;
	std::vector<uint8_t> idat, palette, trns;
	bool have_ihdr = false, done = false;
	for (size_t pos = 8; pos + 12 <= file.size() && !done;) {
		const uint32_t len = be32(&file[pos]);
		const uint8_t* type = &file[pos + 4];
		const uint8_t* data = &file[pos + 8];
		if (pos + 12 + (size_t)len > file.size()) fail("truncated chunk");
		if (!memcmp(type, "IHDR", 4)) {
			if (len != 13) fail("bad IHDR");
			w = be32(data); h = be32(data + 4);
			depth = data[8]; color = data[9]; interlace = data[12];
			if (data[10] != 0 || data[11] != 0) fail("unknown compression / filter method");
			have_ihdr = true;
		} else if (!memcmp(type, "PLTE", 4)) palette.assign(data, data + len);
		else if (!memcmp(type, "tRNS", 4)) trns.assign(data, data + len);
		else if (!memcmp(type, "IDAT", 4)) idat.insert(idat.end(), data, data + len);
		else if (!memcmp(type, "IEND", 4)) done = true;
		pos += 12 + (size_t)len;
	}
	if (!have_ihdr || w == 0 || h == 0 || w > (1u << 16) || h > (1u << 16)) fail("missing or implausible header");
	if (interlace != 0 && interlace != 1) fail("unknown interlace method");
	int channels = 0;
	switch (color) {
		case 0: channels = 1; break;
		case 2: channels = 3; break;
		case 3: channels = 1; break;
		case 4: channels = 2; break;
		case 6: channels = 4; break;
		default: fail("unknown colour type");
	}
	if (!(depth == 8 || depth == 16 || ((color == 0 || color == 3) && (depth == 1 || depth == 2 || depth == 4)))) fail("unsupported bit depth");
	if (color == 3 && (depth == 16 || palette.size() < 3)) fail("bad palette image");

	const size_t bits_per_pixel = (size_t)channels * depth;
	const size_t stride = (w * bits_per_pixel + 7) / 8;
	const size_t bpp = std::max<size_t>(1, bits_per_pixel / 8);   // filter distance in bytes
	// Adam7: seven sub-images (pass p holds the pixels x = x0 + i * dx, y = y0 + j * dy), each filtered on its own
	static const uint32_t kX0[7] = {0, 4, 0, 2, 0, 1, 0}, kY0[7] = {0, 0, 4, 0, 2, 0, 1}, kDx[7] = {8, 8, 4, 4, 2, 2, 1}, kDy[7] = {8, 8, 8, 4, 4, 2, 2};
	struct Pass { uint32_t w, h, x0, y0, dx, dy; size_t stride; };
	std::vector<Pass> passes;
	if (interlace == 0) passes.push_back({w, h, 0, 0, 1, 1, stride});
	else for (int p = 0; p < 7; ++p) {
		const uint32_t pw = (w + kDx[p] - 1 - kX0[p]) / kDx[p], ph = (h + kDy[p] - 1 - kY0[p]) / kDy[p];
		if (kX0[p] < w && kY0[p] < h && pw && ph) passes.push_back({pw, ph, kX0[p], kY0[p], kDx[p], kDy[p], (pw * bits_per_pixel + 7) / 8});
	}
	size_t raw_bytes = 0;
	for (const Pass& p : passes) raw_bytes += (p.stride + 1) * p.h;
	std::vector<uint8_t> raw(raw_bytes);
	{
		z_stream zs{};
		if (inflateInit(&zs) != Z_OK) fail("zlib init failed");
		zs.next_in = idat.data(); zs.avail_in = (uInt)idat.size();
		zs.next_out = raw.data(); zs.avail_out = (uInt)raw.size();
		const int rc = inflate(&zs, Z_FINISH);
		const size_t got = raw.size() - zs.avail_out;
		inflateEnd(&zs);
		if ((rc != Z_STREAM_END && rc != Z_OK && rc != Z_BUF_ERROR) || got != raw.size()) fail("corrupt or truncated image data");
	}

	// unfilter (each scanline is preceded by its filter type); interlaced passes are scattered into the full-size packed image
	std::vector<uint8_t> pix(stride * h, 0);
	size_t raw_at = 0;
	std::vector<uint8_t> line, prev;
	for (const Pass& p : passes) {
		prev.assign(p.stride, 0);
		line.assign(p.stride, 0);
		for (uint32_t j = 0; j < p.h; ++j) {
			const uint8_t ft = raw[raw_at];
			const uint8_t* src = &raw[raw_at + 1];
			raw_at += p.stride + 1;
			uint8_t* dst = interlace == 0 ? &pix[stride * j] : line.data();
			for (size_t i = 0; i < p.stride; ++i) {
				const int a = i >= bpp ? dst[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
				int v = src[i];
				switch (ft) {
					case 0: break;
					case 1: v += a; break;
					case 2: v += b; break;
					case 3: v += (a + b) / 2; break;
					case 4: v += paeth(a, b, c); break;
					default: fail("unknown filter type");
				}
				dst[i] = (uint8_t)v;
			}
			memcpy(prev.data(), dst, p.stride);
			if (interlace == 0) continue;
			uint8_t* out_row = &pix[stride * ((size_t)p.y0 + (size_t)j * p.dy)];
			for (uint32_t i = 0; i < p.w; ++i) {
				const size_t x = (size_t)p.x0 + (size_t)i * p.dx;
				if (bits_per_pixel >= 8) {
					memcpy(out_row + x * (bits_per_pixel / 8), line.data() + i * (bits_per_pixel / 8), bits_per_pixel / 8);
				} else {
					const size_t from = i * bits_per_pixel, to = x * bits_per_pixel;
					const uint32_t v = (line[from / 8] >> (8 - bits_per_pixel - (from % 8))) & ((1u << bits_per_pixel) - 1u);
					out_row[to / 8] |= (uint8_t)(v << (8 - bits_per_pixel - (to % 8)));
				}
			}
		}
	}

	HostImage img;
	img.width = (int)w; img.height = (int)h;
	img.rgba.resize((size_t)w * h * 4);
	auto sample = [&](const uint8_t* row, size_t index) -> uint32_t {   // sample `index` of a row, as stored
		if (depth == 8) return row[index];
		if (depth == 16) return row[2 * index];                         // high byte
		const size_t bit = index * depth;
		return (row[bit / 8] >> (8 - depth - (bit % 8))) & ((1u << depth) - 1u);
	};
	const uint32_t grey_scale = depth < 8 ? 255u / ((1u << depth) - 1u) : 1u;
	for (uint32_t y = 0; y < h; ++y) {
		const uint8_t* row = &pix[stride * y];
		uint8_t* out = &img.rgba[(size_t)y * w * 4];
		for (uint32_t x = 0; x < w; ++x) {
			uint8_t r, g, b, a = 255;
			switch (color) {
				case 0: {
					const uint32_t v = sample(row, x);
					r = g = b = (uint8_t)(v * grey_scale);
					if (trns.size() >= 2 && depth <= 8 && v == (uint32_t)trns[1]) a = 0;
					if (trns.size() >= 2 && depth == 16 && row[2 * x] == trns[0] && row[2 * x + 1] == trns[1]) a = 0;
				} break;
				case 2: {
					r = (uint8_t)sample(row, 3 * x); g = (uint8_t)sample(row, 3 * x + 1); b = (uint8_t)sample(row, 3 * x + 2);
					if (trns.size() >= 6 && depth == 8 && r == trns[1] && g == trns[3] && b == trns[5]) a = 0;
				} break;
				case 3: {
					const uint32_t v = sample(row, x);
					if (3 * v + 2 >= palette.size()) { r = g = b = 0; }
					else { r = palette[3 * v]; g = palette[3 * v + 1]; b = palette[3 * v + 2]; }
					if (v < trns.size()) a = trns[v];
				} break;
				case 4: r = g = b = (uint8_t)sample(row, 2 * x); a = (uint8_t)sample(row, 2 * x + 1); break;
				default: r = (uint8_t)sample(row, 4 * x); g = (uint8_t)sample(row, 4 * x + 1); b = (uint8_t)sample(row, 4 * x + 2); a = (uint8_t)sample(row, 4 * x + 3); break;
			}
			out[4 * x] = r; out[4 * x + 1] = g; out[4 * x + 2] = b; out[4 * x + 3] = a;
		}
	}
	return img;
}

} // namespace ngpb
