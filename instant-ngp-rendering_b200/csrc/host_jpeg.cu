// JPEG decoder on the host whose texels are the reference's. The reference decodes dataset JPEGs with its vendored stb_image
// (stbi_load(..., 4) behind load_stbi, src/common_host.cu:214-217, src/nerf_loader.cu:571): Huffman baseline / extended /
// progressive, 8 bits, 1 / 3 / 4 components. JPEG leaves the arithmetic after entropy decoding to the decoder, so "the same
// image" is only the same bytes when these stages are done the way stb_image does them:
//   dequantisation into 16-bit words; the jidctint-style integer IDCT with 12-bit constants and a 2-bit guard between the
//   passes; chroma upsampling by the 3:1 triangle filters (h2, v2, h2v2) or replication (other ratios); YCbCr -> RGB in
//   20-bit fixed point with the green cross term masked to 16 bits; CMYK / YCCK through the Adobe marker.
// Entropy decoding itself is the standard's and has one result for a valid stream. Corrupt streams fail here where
// stb_image may return a partly decoded picture.
// The loader's default JPEG path stays nvjpeg on the device (host_dataset.cu); this decoder serves ngpb_host_decode_image and
// NGPB_JPEG_DECODER=host, and tests/test_host_io_cpu.py holds it to the reference's stb_image byte for byte.
#include "host_io.h"

#include <algorithm>
#include <cstring>
#include <stdexcept>

namespace ngpb {

namespace {
const uint8_t kDezigzag[64] = {0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
	35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct HuffmanTable {
	bool defined = false;
	int first_code[17] = {}, first_index[17] = {}, count[17] = {};
	uint8_t values[256] = {};
	uint16_t fast[512] = {};   // the next 9 bits -> (code length << 8) | symbol for codes of up to 9 bits, 0 otherwise
};

struct Component {
	int id = 0, h = 1, v = 1, tq = 0, hd = 0, ha = 0, dc_pred = 0;
	int x = 0, y = 0, w2 = 0, h2 = 0;            // true size in samples, padded size in samples
	std::vector<int16_t> coeff;                    // [h2 / 8][w2 / 8][64], natural (row-major) order inside a block
	std::vector<uint8_t> data;                     // [h2][w2] after the IDCT
};

struct Decoder {
	const std::vector<uint8_t>& f;
	const std::string& name;
	size_t at = 2;
	[[noreturn]] void fail(const std::string& why) const { throw std::runtime_error{"Could not open image file '" + name + "': " + why}; }

	int width = 0, height = 0, n_comp = 0, h_max = 1, v_max = 1, mcu_x = 0, mcu_y = 0;
	bool progressive = false, jfif = false, have_frame = false;
	int adobe_transform = -1, rgb_ids = 0, restart_interval = 0;
	uint16_t dequant[4][64] = {};
	HuffmanTable dc_tables[4], ac_tables[4];
	Component comp[4];

	// scan state
	int scan_n = 0, order[4] = {}, spec_start = 0, spec_end = 63, succ_high = 0, succ_low = 0, eob_run = 0;
	uint32_t bit_buffer = 0;
	int bit_count = 0;
	bool hit_marker = false;

	int get8() { if (at >= f.size()) fail("truncated JPEG"); return f[at++]; }
	int get16() { const int hi = get8(); return (hi << 8) | get8(); }

	// ---- bits ----
	void fill() {
		while (bit_count <= 24) {
			uint32_t b = 0;
			if (!hit_marker && at < f.size()) {
				const size_t here = at;
				b = f[at++];
				if (b == 0xFF) {
					size_t q = at;
					int c = q < f.size() ? f[q++] : 0xD9;
					while (c == 0xFF && q < f.size()) c = f[q++];
					if (c != 0) { at = here; hit_marker = true; b = 0; }   // a marker: left where it is for the segment loop
					else at = q;                                            // FF 00 is a data byte of FF
				}
			}
			bit_buffer |= b << (24 - bit_count);
			bit_count += 8;
		}
	}
	int receive(int n) {
		if (n == 0) return 0;
		if (bit_count < n) fill();
		const int v = (int)(bit_buffer >> (32 - n));
		bit_buffer <<= n;
		bit_count -= n;
		return v;
	}
	int receive_extend(int n) {
		if (n == 0) return 0;
		const int v = receive(n);
		return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
	}
	int decode_symbol(const HuffmanTable& t) {
		if (!t.defined) fail("missing Huffman table");
		if (bit_count < 16) fill();
		if (const uint16_t hit = t.fast[bit_buffer >> 23]) {
			bit_buffer <<= hit >> 8;
			bit_count -= hit >> 8;
			return hit & 0xFF;
		}
		int code = 0;
		for (int len = 1; len <= 16; ++len) {
			code = (code << 1) | (int)((bit_buffer >> (32 - len)) & 1u);
			if (t.count[len] && code - t.first_code[len] < t.count[len] && code >= t.first_code[len]) {
				bit_buffer <<= len;
				bit_count -= len;
				return t.values[t.first_index[len] + code - t.first_code[len]];
			}
		}
		fail("bad huffman code");
	}
	void reset_scan_state() {
		bit_buffer = 0; bit_count = 0; hit_marker = false; eob_run = 0;
		for (Component& c : comp) c.dc_pred = 0;
	}

	// ---- segments ----
	void read_quant_tables() {
		int len = get16() - 2;
		while (len > 0) {
			const int q = get8(), precision = q >> 4, t = q & 15;
			if (precision > 1) fail("bad DQT type");
			if (t > 3) fail("bad DQT table");
			for (int i = 0; i < 64; ++i) dequant[t][kDezigzag[i]] = (uint16_t)(precision ? get16() : get8());
			len -= precision ? 129 : 65;
		}
		if (len != 0) fail("bad DQT len");
	}
	void read_huffman_tables() {
		int len = get16() - 2;
		while (len > 0) {
			const int q = get8(), tc = q >> 4, th = q & 15;
			if (tc > 1 || th > 3) fail("bad DHT header");
			HuffmanTable& t = tc == 0 ? dc_tables[th] : ac_tables[th];
			int total = 0;
			for (int i = 1; i <= 16; ++i) { t.count[i] = get8(); total += t.count[i]; }
			if (total > 256) fail("bad DHT header");
			int code = 0, index = 0;
			for (int i = 1; i <= 16; ++i) {
				t.first_code[i] = code; t.first_index[i] = index;
				code += t.count[i]; index += t.count[i];
				if (code > (1 << i)) fail("bad code lengths");
				code <<= 1;
			}
			for (int i = 0; i < total; ++i) t.values[i] = (uint8_t)get8();
			memset(t.fast, 0, sizeof(t.fast));
			for (int len = 1; len <= 9; ++len) for (int k = 0; k < t.count[len]; ++k) {
				const int code_k = t.first_code[len] + k;
				for (int pad = 0; pad < (1 << (9 - len)); ++pad) t.fast[(code_k << (9 - len)) | pad] = (uint16_t)((len << 8) | t.values[t.first_index[len] + k]);
			}
			t.defined = true;
			len -= 17 + total;
		}
		if (len != 0) fail("bad DHT len");
	}
	void read_app(int marker) {
		int len = get16();
		if (len < 2) fail("bad APP len");
		len -= 2;
		const size_t end = at + (size_t)len;
		if (end > f.size()) fail("truncated JPEG");
		if (marker == 0xE0 && len >= 5 && memcmp(&f[at], "JFIF\0", 5) == 0) jfif = true;
		if (marker == 0xEE && len >= 12 && memcmp(&f[at], "Adobe\0", 6) == 0) adobe_transform = f[at + 11];
		at = end;
	}
	void read_frame(int marker) {
		if (have_frame) fail("multiple SOFs");
		progressive = marker == 0xC2;
		const int len = get16();
		if (len < 11) fail("bad SOF len");
		if (get8() != 8) fail("only 8-bit");
		height = get16(); width = get16();
		if (height == 0) fail("no header height");
		if (width == 0) fail("0 width");
		n_comp = get8();
		if (n_comp != 1 && n_comp != 3 && n_comp != 4) fail("bad component count");
		if (len != 8 + 3 * n_comp) fail("bad SOF len");
		static const char rgb[3] = {'R', 'G', 'B'};
		for (int i = 0; i < n_comp; ++i) {
			Component& c = comp[i];
			c.id = get8();
			if (n_comp == 3 && c.id == rgb[i]) ++rgb_ids;
			const int q = get8();
			c.h = q >> 4; c.v = q & 15; c.tq = get8();
			if (c.h < 1 || c.h > 4) fail("bad H");
			if (c.v < 1 || c.v > 4) fail("bad V");
			if (c.tq > 3) fail("bad TQ");
			h_max = std::max(h_max, c.h); v_max = std::max(v_max, c.v);
		}
		for (int i = 0; i < n_comp; ++i) if (h_max % comp[i].h || v_max % comp[i].v) fail("bad H");
		mcu_x = (width + h_max * 8 - 1) / (h_max * 8);
		mcu_y = (height + v_max * 8 - 1) / (v_max * 8);
		if ((size_t)width * height > f.size() * 4096) fail("JPEG size implausible for the file");
		for (int i = 0; i < n_comp; ++i) {
			Component& c = comp[i];
			c.x = (width * c.h + h_max - 1) / h_max;
			c.y = (height * c.v + v_max - 1) / v_max;
			c.w2 = mcu_x * c.h * 8; c.h2 = mcu_y * c.v * 8;
			c.coeff.assign((size_t)c.w2 * c.h2, 0);
			c.data.assign((size_t)c.w2 * c.h2, 0);
		}
		have_frame = true;
	}
	void read_scan_header() {
		if (!have_frame) fail("SOS before SOF");
		const int len = get16();
		scan_n = get8();
		if (scan_n < 1 || scan_n > 4 || scan_n > n_comp) fail("bad SOS component count");
		if (len != 6 + 2 * scan_n) fail("bad SOS len");
		for (int i = 0; i < scan_n; ++i) {
			const int id = get8(), q = get8();
			int which = 0;
			while (which < n_comp && comp[which].id != id) ++which;
			if (which == n_comp) fail("bad component id in SOS");
			comp[which].hd = q >> 4; comp[which].ha = q & 15;
			if (comp[which].hd > 3 || comp[which].ha > 3) fail("bad huffman table index");
			order[i] = which;
		}
		spec_start = get8(); spec_end = get8();
		const int aa = get8();
		succ_high = aa >> 4; succ_low = aa & 15;
		if (progressive) {
			if (spec_start > 63 || spec_end > 63 || spec_start > spec_end || succ_high > 13 || succ_low > 13) fail("bad SOS");
		} else {
			if (spec_start != 0 || succ_high != 0 || succ_low != 0) fail("bad SOS");
			spec_end = 63;
		}
	}

	// ---- blocks ----
	void block_baseline(Component& c, int16_t* d) {
		const int t = decode_symbol(dc_tables[c.hd]);
		if (t > 15) fail("bad huffman code");
		c.dc_pred += receive_extend(t);
		d[0] = (int16_t)c.dc_pred;
		const HuffmanTable& ac = ac_tables[c.ha];
		for (int k = 1; k < 64;) {
			const int rs = decode_symbol(ac), s = rs & 15, r = rs >> 4;
			if (s == 0) {
				if (rs != 0xF0) break;
				k += 16;
			} else {
				k += r;
				if (k > 63) fail("bad huffman code");
				d[kDezigzag[k++]] = (int16_t)receive_extend(s);
			}
		}
	}
	void block_progressive_dc(Component& c, int16_t* d) {
		if (spec_end != 0) fail("can't merge dc and ac");
		if (succ_high == 0) {
			const int t = decode_symbol(dc_tables[c.hd]);
			if (t > 15) fail("bad huffman code");
			c.dc_pred += receive_extend(t);
			d[0] = (int16_t)(c.dc_pred * (1 << succ_low));
		} else if (receive(1)) {
			d[0] = (int16_t)(d[0] + (int16_t)(1 << succ_low));
		}
	}
	void refine(int16_t& p, int16_t bit) {
		if (receive(1) && (p & bit) == 0) p = (int16_t)(p > 0 ? p + bit : p - bit);
	}
	void block_progressive_ac(Component& c, int16_t* d) {
		if (spec_start == 0) fail("can't merge dc and ac");
		const HuffmanTable& ac = ac_tables[c.ha];
		if (succ_high == 0) {
			if (eob_run) { --eob_run; return; }
			int k = spec_start;
			do {
				const int rs = decode_symbol(ac), s = rs & 15, r = rs >> 4;
				if (s == 0) {
					if (r < 15) {
						eob_run = (1 << r) + (r ? receive(r) : 0) - 1;
						break;
					}
					k += 16;
				} else {
					k += r;
					if (k > 63) fail("bad huffman code");
					d[kDezigzag[k++]] = (int16_t)(receive_extend(s) * (1 << succ_low));
				}
			} while (k <= spec_end);
			return;
		}
		const int16_t bit = (int16_t)(1 << succ_low);
		if (eob_run) {
			--eob_run;
			for (int k = spec_start; k <= spec_end; ++k) {
				int16_t& p = d[kDezigzag[k]];
				if (p != 0) refine(p, bit);
			}
			return;
		}
		int k = spec_start;
		do {
			const int rs = decode_symbol(ac);
			int s = rs & 15, r = rs >> 4;
			if (s == 0) {
				if (r < 15) {
					eob_run = (1 << r) - 1 + (r ? receive(r) : 0);
					r = 64;   // to the end of the band
				}
			} else {
				if (s != 1) fail("bad huffman code");
				s = receive(1) ? bit : -bit;
			}
			while (k <= spec_end) {
				int16_t& p = d[kDezigzag[k++]];
				if (p != 0) {
					refine(p, bit);
				} else {
					if (r == 0) { p = (int16_t)s; break; }
					--r;
				}
			}
		} while (k <= spec_end);
	}
	void decode_block(Component& c, int bx, int by) {
		int16_t* d = &c.coeff[((size_t)by * (c.w2 / 8) + bx) * 64];
		if (!progressive) block_baseline(c, d);
		else if (spec_start == 0) block_progressive_dc(c, d);
		else block_progressive_ac(c, d);
	}
	// after each MCU: a restart interval ends with an RSTn marker and resets the predictors
	bool mcu_done(int& todo) {
		if (restart_interval == 0 || --todo > 0) return true;
		if (bit_count < 24) fill();
		if (at + 1 < f.size() && f[at] == 0xFF && f[at + 1] >= 0xD0 && f[at + 1] <= 0xD7) {
			at += 2;
			reset_scan_state();
			todo = restart_interval;
			return true;
		}
		return false;   // no restart marker where one is due: stb_image stops the scan here and keeps what it has
	}
	void read_scan() {
		read_scan_header();
		reset_scan_state();
		int todo = restart_interval;
		if (scan_n == 1) {
			Component& c = comp[order[0]];
			const int w = (c.x + 7) >> 3, h = (c.y + 7) >> 3;
			for (int j = 0; j < h; ++j) for (int i = 0; i < w; ++i) {
				decode_block(c, i, j);
				if (!mcu_done(todo)) return;
			}
		} else {
			if (progressive && spec_start != 0) fail("can't merge dc and ac");
			for (int j = 0; j < mcu_y; ++j) for (int i = 0; i < mcu_x; ++i) {
				for (int k = 0; k < scan_n; ++k) {
					Component& c = comp[order[k]];
					for (int y = 0; y < c.v; ++y) for (int x = 0; x < c.h; ++x) decode_block(c, i * c.h + x, j * c.v + y);
				}
				if (!mcu_done(todo)) return;
			}
		}
	}

	// ---- reconstruction ----
	static uint8_t clamp(int x) { return (uint8_t)((unsigned)x > 255 ? (x < 0 ? 0 : 255) : x); }
	static int fixed12(double x) { return (int)(x * 4096 + 0.5); }
	// one 8-point pass of the jidctint "slow" transform, constants in 12-bit fixed point. Sums and products wrap modulo 2^32
	// (unsigned arithmetic, read back as signed): what the reference's build does on the oversized coefficients of a corrupt
	// stream, without the undefined behaviour of signed overflow.
	typedef uint32_t U;
	static U mul(U a, int c) { return a * (U)c; }
	static void idct_1d(const int s[8], U& x0, U& x1, U& x2, U& x3, U& t0, U& t1, U& t2, U& t3) {
		U p2 = (U)s[2], p3 = (U)s[6];
		U p1 = mul(p2 + p3, fixed12(0.5411961f));
		t2 = p1 + mul(p3, fixed12(-1.847759065f));
		t3 = p1 + mul(p2, fixed12(0.765366865f));
		p2 = (U)s[0]; p3 = (U)s[4];
		t0 = mul(p2 + p3, 4096);
		t1 = mul(p2 - p3, 4096);
		x0 = t0 + t3; x3 = t0 - t3; x1 = t1 + t2; x2 = t1 - t2;
		t0 = (U)s[7]; t1 = (U)s[5]; t2 = (U)s[3]; t3 = (U)s[1];
		p3 = t0 + t2;
		U p4 = t1 + t3;
		p1 = t0 + t3; p2 = t1 + t2;
		const U p5 = mul(p3 + p4, fixed12(1.175875602f));
		t0 = mul(t0, fixed12(0.298631336f));
		t1 = mul(t1, fixed12(2.053119869f));
		t2 = mul(t2, fixed12(3.072711026f));
		t3 = mul(t3, fixed12(1.501321110f));
		p1 = p5 + mul(p1, fixed12(-0.899976223f));
		p2 = p5 + mul(p2, fixed12(-2.562915447f));
		p3 = mul(p3, fixed12(-1.961570560f));
		p4 = mul(p4, fixed12(-0.390180644f));
		t3 += p1 + p4; t2 += p2 + p3; t1 += p2 + p4; t0 += p1 + p3;
	}
	static int sar(U x, int n) { return (int)((int32_t)x >> n); }
	static void idct_block(uint8_t* out, int stride, const int16_t d[64]) {
		int val[64];
		for (int i = 0; i < 8; ++i) {
			int* v = val + i;
			if (!d[i + 8] && !d[i + 16] && !d[i + 24] && !d[i + 32] && !d[i + 40] && !d[i + 48] && !d[i + 56]) {
				const int dc = d[i] * 4;
				for (int r = 0; r < 8; ++r) v[r * 8] = dc;
				continue;
			}
			const int s[8] = {d[i], d[i + 8], d[i + 16], d[i + 24], d[i + 32], d[i + 40], d[i + 48], d[i + 56]};
			U x0, x1, x2, x3, t0, t1, t2, t3;
			idct_1d(s, x0, x1, x2, x3, t0, t1, t2, t3);
			x0 += 512; x1 += 512; x2 += 512; x3 += 512;
			v[0] = sar(x0 + t3, 10); v[56] = sar(x0 - t3, 10);
			v[8] = sar(x1 + t2, 10); v[48] = sar(x1 - t2, 10);
			v[16] = sar(x2 + t1, 10); v[40] = sar(x2 - t1, 10);
			v[24] = sar(x3 + t0, 10); v[32] = sar(x3 - t0, 10);
		}
		for (int i = 0; i < 8; ++i, out += stride) {
			U x0, x1, x2, x3, t0, t1, t2, t3;
			idct_1d(val + i * 8, x0, x1, x2, x3, t0, t1, t2, t3);
			const U bias = 65536u + (128u << 17);
			x0 += bias; x1 += bias; x2 += bias; x3 += bias;
			out[0] = clamp(sar(x0 + t3, 17)); out[7] = clamp(sar(x0 - t3, 17));
			out[1] = clamp(sar(x1 + t2, 17)); out[6] = clamp(sar(x1 - t2, 17));
			out[2] = clamp(sar(x2 + t1, 17)); out[5] = clamp(sar(x2 - t1, 17));
			out[3] = clamp(sar(x3 + t0, 17)); out[4] = clamp(sar(x3 - t0, 17));
		}
	}
	void reconstruct() {
		for (int n = 0; n < n_comp; ++n) {
			Component& c = comp[n];
			const uint16_t* q = dequant[c.tq];
			const int bw = c.w2 / 8, bh = c.h2 / 8;
			for (int by = 0; by < bh; ++by) for (int bx = 0; bx < bw; ++bx) {
				int16_t* d = &c.coeff[((size_t)by * bw + bx) * 64];
				for (int i = 0; i < 64; ++i) d[i] = (int16_t)(d[i] * q[i]);
				idct_block(&c.data[(size_t)by * 8 * c.w2 + (size_t)bx * 8], c.w2, d);
			}
		}
	}

	// one output row of a component at full width: the row itself, the 3:1 triangle filters, or replication
	const uint8_t* upsample_row(const Component& c, const uint8_t* near_row, const uint8_t* far_row, std::vector<uint8_t>& line) const {
		const int hs = h_max / c.h, vs = v_max / c.v, w = (width + hs - 1) / hs;
		uint8_t* out = line.data();
		if (hs == 1 && vs == 1) return near_row;
		if (hs == 1 && vs == 2) {
			for (int i = 0; i < w; ++i) out[i] = (uint8_t)((3 * near_row[i] + far_row[i] + 2) >> 2);
		} else if (hs == 2 && vs == 1) {
			if (w == 1) { out[0] = out[1] = near_row[0]; return out; }
			out[0] = near_row[0];
			out[1] = (uint8_t)((near_row[0] * 3 + near_row[1] + 2) >> 2);
			int i = 1;
			for (; i < w - 1; ++i) {
				const int n = 3 * near_row[i] + 2;
				out[i * 2] = (uint8_t)((n + near_row[i - 1]) >> 2);
				out[i * 2 + 1] = (uint8_t)((n + near_row[i + 1]) >> 2);
			}
			out[i * 2] = (uint8_t)((near_row[w - 2] * 3 + near_row[w - 1] + 2) >> 2);
			out[i * 2 + 1] = near_row[w - 1];
		} else if (hs == 2 && vs == 2) {
			if (w == 1) { out[0] = out[1] = (uint8_t)((3 * near_row[0] + far_row[0] + 2) >> 2); return out; }
			int t1 = 3 * near_row[0] + far_row[0];
			out[0] = (uint8_t)((t1 + 2) >> 2);
			for (int i = 1; i < w; ++i) {
				const int t0 = t1;
				t1 = 3 * near_row[i] + far_row[i];
				out[i * 2 - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4);
				out[i * 2] = (uint8_t)((3 * t1 + t0 + 8) >> 4);
			}
			out[w * 2 - 1] = (uint8_t)((t1 + 2) >> 2);
		} else {
			for (int i = 0; i < w; ++i) for (int j = 0; j < hs; ++j) out[i * hs + j] = near_row[i];
		}
		return out;
	}
	static uint8_t blinn(uint8_t x, uint8_t y) {
		const unsigned t = x * y + 128;
		return (uint8_t)((t + (t >> 8)) >> 8);
	}
	static int ycc_fixed(float x) { return ((int)(x * 4096.0f + 0.5f)) << 8; }
	static void ycc_to_rgb(uint8_t* out, const uint8_t* y, const uint8_t* pcb, const uint8_t* pcr, int count) {
		for (int i = 0; i < count; ++i, out += 4) {
			const int y_fixed = (y[i] << 20) + (1 << 19);
			const int cr = pcr[i] - 128, cb = pcb[i] - 128;
			int r = y_fixed + cr * ycc_fixed(1.40200f);
			int g = y_fixed + (cr * -ycc_fixed(0.71414f)) + (int)((unsigned)(cb * -ycc_fixed(0.34414f)) & 0xffff0000u);
			int b = y_fixed + cb * ycc_fixed(1.77200f);
			out[0] = clamp(r >> 20); out[1] = clamp(g >> 20); out[2] = clamp(b >> 20); out[3] = 255;
		}
	}
	HostImage assemble() {
		HostImage img;
		img.width = width; img.height = height;
		img.rgba.resize((size_t)width * height * 4);
		const bool is_rgb = n_comp == 3 && (rgb_ids == 3 || (adobe_transform == 0 && !jfif));
		struct Rows { int ystep, ypos; const uint8_t* line0; const uint8_t* line1; };
		Rows rows[4];
		std::vector<uint8_t> lines[4];
		for (int k = 0; k < n_comp; ++k) {
			rows[k] = {(v_max / comp[k].v) >> 1, 0, comp[k].data.data(), comp[k].data.data()};
			lines[k].resize((size_t)width + 8);
		}
		const uint8_t* cout[4] = {};
		for (int j = 0; j < height; ++j) {
			uint8_t* out = &img.rgba[(size_t)j * width * 4];
			for (int k = 0; k < n_comp; ++k) {
				Rows& r = rows[k];
				const int vs = v_max / comp[k].v;
				const bool y_bot = r.ystep >= (vs >> 1);
				cout[k] = upsample_row(comp[k], y_bot ? r.line1 : r.line0, y_bot ? r.line0 : r.line1, lines[k]);
				if (++r.ystep >= vs) {
					r.ystep = 0;
					r.line0 = r.line1;
					if (++r.ypos < comp[k].y) r.line1 += comp[k].w2;
				}
			}
			if (n_comp == 3) {
				if (is_rgb) for (int i = 0; i < width; ++i, out += 4) { out[0] = cout[0][i]; out[1] = cout[1][i]; out[2] = cout[2][i]; out[3] = 255; }
				else ycc_to_rgb(out, cout[0], cout[1], cout[2], width);
			} else if (n_comp == 4) {
				if (adobe_transform == 0) {
					for (int i = 0; i < width; ++i, out += 4) {
						const uint8_t m = cout[3][i];
						out[0] = blinn(cout[0][i], m); out[1] = blinn(cout[1][i], m); out[2] = blinn(cout[2][i], m); out[3] = 255;
					}
				} else if (adobe_transform == 2) {
					ycc_to_rgb(out, cout[0], cout[1], cout[2], width);
					for (int i = 0; i < width; ++i, out += 4) {
						const uint8_t m = cout[3][i];
						out[0] = blinn((uint8_t)(255 - out[0]), m); out[1] = blinn((uint8_t)(255 - out[1]), m); out[2] = blinn((uint8_t)(255 - out[2]), m);
					}
				} else {
					ycc_to_rgb(out, cout[0], cout[1], cout[2], width);
				}
			} else {
				for (int i = 0; i < width; ++i, out += 4) { out[0] = out[1] = out[2] = cout[0][i]; out[3] = 255; }
			}
		}
		return img;
	}

	HostImage run() {
		if (f.size() < 4 || f[0] != 0xFF || f[1] != 0xD8) fail("no SOI");
		bool seen_scan = false;
		for (;;) {
			// next marker: skip anything that is not 0xFF, then fill bytes
			while (at < f.size() && f[at] != 0xFF) ++at;
			while (at < f.size() && f[at] == 0xFF) ++at;
			if (at >= f.size()) break;
			const int m = f[at++];
			if (m == 0xD9) break;
			if (m == 0x00 || (m >= 0xD0 && m <= 0xD7) || m == 0x01) continue;
			if (m == 0xC0 || m == 0xC1 || m == 0xC2) read_frame(m);
			else if (m == 0xC4) read_huffman_tables();
			else if (m == 0xDB) read_quant_tables();
			else if (m == 0xDD) { if (get16() != 4) fail("bad DRI len"); restart_interval = get16(); }
			else if (m == 0xDA) { read_scan(); seen_scan = true; }
			else if ((m >= 0xE0 && m <= 0xEF) || m == 0xFE) read_app(m);
			else if (m == 0xDC) { if (get16() != 4) fail("bad DNL len"); if (get16() != height) fail("bad DNL height"); }
			else fail("unknown marker");
		}
		if (!have_frame || !seen_scan) fail("no image in the JPEG stream");
		reconstruct();
		return assemble();
	}
};
} // namespace

HostImage decode_jpeg(const std::vector<uint8_t>& file, const std::string& name) {
	Decoder d{file, name};
	return d.run();
}

} // namespace ngpb
