// drt_jpeg.cpp -- JPEG decoder (baseline and progressive Huffman, 8 bit) for scene textures.
//
// The reference decodes embedded images with stb_image 2.29 (vendored: DustRayTracer/thirdparty/stb_image/stb_image.h, called
// from DustRayTracer/src/Core/Scene/Texture.cu:21-30 with req_comp = 0; call site Scene.cu:88-117). Texels feed the albedo
// (Texture.cu:33-58), so the decoder has to produce stb's BYTES, not just "a" JPEG decode: every step that rounds is restated
// with stb's arithmetic, each citing the stb_image.h lines it follows. Entropy decoding is the T.81 procedure (Annex F / G),
// which has one right answer; tables and bit reader are ours.
//   inverse DCT       stb_image.h:2425-2521  (IJG "islow" in 12-bit fixed point, 2 extra bits kept after the column pass)
//   dequantisation    stb_image.h:2228,2251,3072-3077 (16-bit products)
//   chroma upsampling stb_image.h:3464-3657  (JFIF-centred triangle filters: v2, h2, hv2; nearest for other ratios)
//   YCbCr -> RGB      stb_image.h:3659-3683  (20-bit fixed point with the constants reduced to 12 bits)
//   colour handling   stb_image.h:3864-4024  (component ids 'R','G','B' or Adobe transform 0 without JFIF = RGB;
//                                             4 components = CMYK / YCCK by the Adobe tag); output 1 or 3 channels
// Pinned by tests/test_jpeg.py: texel CRCs of the reference's own JPEG-textured models decoded by the reference's stb build,
// and a matrix of encoder settings (progressive, 4:4:4 / 4:2:2 / 4:4:0 / 4:2:0 / 4:1:1, grey, restart intervals, odd sizes).
#include "drt_host.h"

#include <cstring>
#include <vector>

namespace drt {

namespace {

const uint8_t kZigzag[64 + 16] = { // T.81 figure A.6; the 16 extra entries absorb runs that overshoot a corrupt block
	0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
	35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
	63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63 };

struct Huff {
	// canonical code, T.81 F.2.2.3: codes of length L are mincode[L] .. maxcode[L], symbol index = valptr[L] + code - mincode[L]
	int maxcode[18], mincode[17], valptr[17];
	uint8_t values[256];
	int count = 0;
	bool defined = false;
};

struct Component {
	int id = 0, h = 1, v = 1, tq = 0, hd = 0, ha = 0;
	int dc_pred = 0;
	int x = 0, y = 0;        // effective samples of this component
	int w2 = 0, h2 = 0;      // allocated plane (whole MCUs)
	int blocks_w = 0;        // coefficient blocks per row (progressive)
	std::vector<uint8_t> plane;
	std::vector<int16_t> coeff;
};

struct Decoder {
	const uint8_t* p = nullptr;
	const uint8_t* end = nullptr;
	// entropy-coded segment reader: 32-bit window, MSB first; a marker stops the feed (zeros follow), like T.81 F.2.2.5
	uint32_t window = 0;
	int bits = 0;
	int marker = 0;          // marker met while feeding (0 = none)
	bool starved = false;
	int width = 0, height = 0, ncomp = 0;
	bool progressive = false;
	bool jfif = false;
	int adobe_transform = -1;
	int rgb_ids = 0;
	int h_max = 1, v_max = 1, mcu_x = 0, mcu_y = 0;
	int restart_interval = 0;
	int scan_n = 0, order[4] = { 0, 0, 0, 0 };
	int ss = 0, se = 63, ah = 0, al = 0;
	int eob_run = 0;
	uint16_t quant[4][64];
	Huff dc[4], ac[4];
	Component comp[4];
	std::string err;

	bool fail(const char* m) { if (err.empty()) err = m; return false; }
	int byte() { return p < end ? *p++ : 0; }
	bool at_end() const { return p >= end; }
	int be16() { int a = byte(); return (a << 8) | byte(); }

	void feed()
	{
		do {
			int b = starved ? 0 : byte();
			if (b == 0xff) {
				int c = byte();
				while (c == 0xff) c = byte(); // fill bytes
				if (c != 0) { marker = c; starved = true; return; }
			}
			window |= (uint32_t)b << (24 - bits);
			bits += 8;
		} while (bits <= 24);
	}
	int take(int n) // n unsigned bits (0 when the segment has run dry)
	{
		if (bits < n) feed();
		if (bits < n) return 0;
		const uint32_t v = n ? window >> (32 - n) : 0u;
		window = n >= 32 ? 0u : window << n;
		bits -= n;
		return (int)v;
	}
	int take_bit() { return take(1); }
	// RECEIVE + EXTEND (T.81 F.2.2.1, figure F.12)
	int receive_extend(int n)
	{
		if (bits < n) feed();
		if (bits < n) return 0;
		const int v = take(n);
		return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v;
	}
	int decode_symbol(const Huff& h)
	{
		if (bits < 16) feed();
		const uint32_t top = window >> 16; // the next 16 bits
		for (int len = 1; len <= 16; len++) {
			const int code = (int)(top >> (16 - len));
			if (h.maxcode[len] >= 0 && code <= h.maxcode[len] && code >= h.mincode[len]) {
				if (len > bits) return -1;
				window <<= len;
				bits -= len;
				const int idx = h.valptr[len] + code - h.mincode[len];
				return idx < h.count ? h.values[idx] : -1;
			}
		}
		return -1;
	}
	void reset_entropy()
	{
		bits = 0; window = 0; starved = false; marker = 0; eob_run = 0;
		for (int i = 0; i < 4; i++) comp[i].dc_pred = 0;
	}
};

bool build_huff(Huff& h, const int counts[16])
{
	int code = 0, k = 0;
	for (int len = 1; len <= 16; len++) {
		h.valptr[len] = k;
		h.mincode[len] = code;
		if (counts[len - 1] > 0) {
			code += counts[len - 1];
			k += counts[len - 1];
			if (code - 1 >= (1 << len)) return false;
			h.maxcode[len] = code - 1;
		}
		else h.maxcode[len] = -1;
		code <<= 1;
	}
	h.count = k;
	h.defined = true;
	return k <= 256;
}

inline uint8_t clamp255(int x) { return (uint8_t)(x < 0 ? 0 : (x > 255 ? 255 : x)); }

// One 1-D pass of the IJG "islow" inverse DCT with constants scaled by 4096 (stb_image.h:2429-2464). Outputs the even part
// x0..x3 and the odd part t0..t3; the caller combines them (x0 +- t3, x1 +- t2, x2 +- t1, x3 +- t0).
struct Idct1D { int x0, x1, x2, x3, t0, t1, t2, t3; };
inline Idct1D idct_1d(int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7)
{
	Idct1D r;
	int p1 = (s2 + s6) * 2217;            // 0.5411961
	const int e2 = p1 + s6 * -7567;       // -1.847759065
	const int e3 = p1 + s2 * 3135;        // 0.765366865
	const int e0 = (s0 + s4) * 4096, e1 = (s0 - s4) * 4096;
	r.x0 = e0 + e3; r.x3 = e0 - e3; r.x1 = e1 + e2; r.x2 = e1 - e2;
	int t0 = s7, t1 = s5, t2 = s3, t3 = s1;
	int p3 = t0 + t2, p4 = t1 + t3;
	p1 = t0 + t3;
	int p2 = t1 + t2;
	const int p5 = (p3 + p4) * 4816;      // 1.175875602
	t0 *= 1223;                           // 0.298631336
	t1 *= 8410;                           // 2.053119869
	t2 *= 12586;                          // 3.072711026
	t3 *= 6149;                           // 1.501321110
	p1 = p5 + p1 * -3685;                 // -0.899976223
	p2 = p5 + p2 * -10497;                // -2.562915447
	p3 *= -8034;                          // -1.961570560
	p4 *= -1597;                          // -0.390180644
	r.t3 = t3 + p1 + p4; r.t2 = t2 + p2 + p3; r.t1 = t1 + p2 + p4; r.t0 = t0 + p1 + p3;
	return r;
}

// 8x8 block: columns first (results keep 2 extra bits: +512 >> 10), then rows (+65536 rounds the 17-bit scale, +128 << 17
// recentres), clamped to 0..255 (stb_image.h:2466-2521).
void idct_block(uint8_t* out, int stride, const int16_t d[64])
{
	int v[64];
	for (int i = 0; i < 8; i++) {
		Idct1D r = idct_1d(d[i], d[8 + i], d[16 + i], d[24 + i], d[32 + i], d[40 + i], d[48 + i], d[56 + i]);
		r.x0 += 512; r.x1 += 512; r.x2 += 512; r.x3 += 512;
		v[i] = (r.x0 + r.t3) >> 10; v[56 + i] = (r.x0 - r.t3) >> 10;
		v[8 + i] = (r.x1 + r.t2) >> 10; v[48 + i] = (r.x1 - r.t2) >> 10;
		v[16 + i] = (r.x2 + r.t1) >> 10; v[40 + i] = (r.x2 - r.t1) >> 10;
		v[24 + i] = (r.x3 + r.t0) >> 10; v[32 + i] = (r.x3 - r.t0) >> 10;
	}
	for (int i = 0; i < 8; i++) {
		const int* w = v + 8 * i;
		uint8_t* o = out + (size_t)stride * i;
		Idct1D r = idct_1d(w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]);
		const int bias = 65536 + (128 << 17);
		r.x0 += bias; r.x1 += bias; r.x2 += bias; r.x3 += bias;
		o[0] = clamp255((r.x0 + r.t3) >> 17); o[7] = clamp255((r.x0 - r.t3) >> 17);
		o[1] = clamp255((r.x1 + r.t2) >> 17); o[6] = clamp255((r.x1 - r.t2) >> 17);
		o[2] = clamp255((r.x2 + r.t1) >> 17); o[5] = clamp255((r.x2 - r.t1) >> 17);
		o[3] = clamp255((r.x3 + r.t0) >> 17); o[4] = clamp255((r.x3 - r.t0) >> 17);
	}
}

// Sequential block (T.81 F.2.2): coefficients are dequantised as they are decoded, products truncated to 16 bits
// (stb_image.h:2209-2262).
bool decode_block_sequential(Decoder& D, int16_t data[64], int ci)
{
	Component& c = D.comp[ci];
	const uint16_t* q = D.quant[c.tq];
	const int t = D.decode_symbol(D.dc[c.hd]);
	if (t < 0 || t > 15) return D.fail("JPEG: bad DC Huffman code");
	memset(data, 0, 64 * sizeof(int16_t));
	const int diff = t ? D.receive_extend(t) : 0;
	c.dc_pred += diff;
	data[0] = (int16_t)(c.dc_pred * q[0]);
	int k = 1;
	do {
		const int rs = D.decode_symbol(D.ac[c.ha]);
		if (rs < 0) return D.fail("JPEG: bad AC Huffman code");
		const int s = rs & 15, r = rs >> 4;
		if (s == 0) {
			if (rs != 0xf0) break; // end of block
			k += 16;
		}
		else {
			k += r;
			const int z = kZigzag[k++];
			data[z] = (int16_t)(D.receive_extend(s) * q[z]);
		}
	} while (k < 64);
	return true;
}

// Progressive DC scan, first pass and refinement (T.81 G.1.2.1; stb_image.h:2264-2290).
bool decode_block_prog_dc(Decoder& D, int16_t data[64], int ci)
{
	if (D.se != 0) return D.fail("JPEG: DC scan with AC coefficients");
	Component& c = D.comp[ci];
	if (D.ah == 0) {
		memset(data, 0, 64 * sizeof(int16_t));
		const int t = D.decode_symbol(D.dc[c.hd]);
		if (t < 0 || t > 15) return D.fail("JPEG: bad DC Huffman code");
		const int diff = t ? D.receive_extend(t) : 0;
		c.dc_pred += diff;
		data[0] = (int16_t)(c.dc_pred * (1 << D.al));
	}
	else if (D.take_bit()) data[0] = (int16_t)(data[0] + (1 << D.al));
	return true;
}

// Progressive AC scan, first pass and refinement (T.81 G.1.2.2, G.1.2.3; stb_image.h:2294-2412).
bool decode_block_prog_ac(Decoder& D, int16_t data[64], int ci)
{
	if (D.ss == 0) return D.fail("JPEG: AC scan starting at the DC coefficient");
	const Huff& h = D.ac[D.comp[ci].ha];
	if (D.ah == 0) {
		if (D.eob_run) { --D.eob_run; return true; }
		int k = D.ss;
		do {
			const int rs = D.decode_symbol(h);
			if (rs < 0) return D.fail("JPEG: bad AC Huffman code");
			const int s = rs & 15, r = rs >> 4;
			if (s == 0) {
				if (r < 15) {
					D.eob_run = 1 << r;
					if (r) D.eob_run += D.take(r);
					--D.eob_run;
					break;
				}
				k += 16;
			}
			else {
				k += r;
				const int z = kZigzag[k++];
				data[z] = (int16_t)(D.receive_extend(s) * (1 << D.al));
			}
		} while (k <= D.se);
		return true;
	}
	const int16_t bit = (int16_t)(1 << D.al);
	auto refine = [&](int16_t& v) { // one correction bit for a coefficient that is already non-zero
		if (D.take_bit() && (v & bit) == 0) v = (int16_t)(v > 0 ? v + bit : v - bit);
	};
	if (D.eob_run) {
		--D.eob_run;
		for (int k = D.ss; k <= D.se; k++) {
			int16_t& v = data[kZigzag[k]];
			if (v != 0) refine(v);
		}
		return true;
	}
	int k = D.ss;
	do {
		const int rs = D.decode_symbol(h);
		if (rs < 0) return D.fail("JPEG: bad AC Huffman code");
		int s = rs & 15, r = rs >> 4;
		if (s == 0) {
			if (r < 15) {
				D.eob_run = (1 << r) - 1;
				if (r) D.eob_run += D.take(r);
				r = 64; // to the end of the band
			}
		}
		else {
			if (s != 1) return D.fail("JPEG: bad refinement code");
			s = D.take_bit() ? bit : -bit;
		}
		while (k <= D.se) {
			int16_t& v = data[kZigzag[k++]];
			if (v != 0) refine(v);
			else {
				if (r == 0) { v = (int16_t)s; break; }
				--r;
			}
		}
	} while (k <= D.se);
	return true;
}

// After each MCU: restart bookkeeping (T.81 E.2.4). Returns false when the segment ends here without a restart marker.
bool mcu_done(Decoder& D, int& todo)
{
	if (--todo > 0) return true;
	if (D.bits < 24) D.feed();
	if (!(D.marker >= 0xd0 && D.marker <= 0xd7)) return false;
	D.reset_entropy();
	todo = D.restart_interval ? D.restart_interval : 0x7fffffff;
	return true;
}

bool decode_scan(Decoder& D)
{
	D.reset_entropy();
	int todo = D.restart_interval ? D.restart_interval : 0x7fffffff;
	int16_t block[64];
	if (D.scan_n == 1) {
		// one component: its blocks in raster order over its own effective size (T.81 A.2.2)
		const int ci = D.order[0];
		Component& c = D.comp[ci];
		const int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3;
		for (int j = 0; j < bh; j++)
			for (int i = 0; i < bw; i++) {
				if (!D.progressive) {
					if (!decode_block_sequential(D, block, ci)) return false;
					idct_block(c.plane.data() + (size_t)c.w2 * j * 8 + i * 8, c.w2, block);
				}
				else {
					int16_t* data = c.coeff.data() + 64 * ((size_t)i + (size_t)j * c.blocks_w);
					if (D.ss == 0 ? !decode_block_prog_dc(D, data, ci) : !decode_block_prog_ac(D, data, ci)) return false;
				}
				if (!mcu_done(D, todo)) return true;
			}
		return true;
	}
	for (int j = 0; j < D.mcu_y; j++)
		for (int i = 0; i < D.mcu_x; i++) {
			for (int k = 0; k < D.scan_n; k++) {
				const int ci = D.order[k];
				Component& c = D.comp[ci];
				for (int y = 0; y < c.v; y++)
					for (int x = 0; x < c.h; x++) {
						const int bx = i * c.h + x, by = j * c.v + y;
						if (!D.progressive) {
							if (!decode_block_sequential(D, block, ci)) return false;
							idct_block(c.plane.data() + (size_t)c.w2 * by * 8 + bx * 8, c.w2, block);
						}
						else if (!decode_block_prog_dc(D, c.coeff.data() + 64 * ((size_t)bx + (size_t)by * c.blocks_w), ci)) return false;
					}
			}
			if (!mcu_done(D, todo)) return true;
		}
	return true;
}

bool read_tables(Decoder& D, int m)
{
	if (m == 0xdd) { // DRI
		if (D.be16() != 4) return D.fail("JPEG: bad DRI length");
		D.restart_interval = D.be16();
		return true;
	}
	if (m == 0xdb) { // DQT
		int L = D.be16() - 2;
		while (L > 0) {
			const int q = D.byte(), wide = q >> 4, t = q & 15;
			if (wide > 1 || t > 3) return D.fail("JPEG: bad DQT");
			for (int i = 0; i < 64; i++) D.quant[t][kZigzag[i]] = (uint16_t)(wide ? D.be16() : D.byte());
			L -= wide ? 129 : 65;
		}
		return L == 0 ? true : D.fail("JPEG: bad DQT length");
	}
	if (m == 0xc4) { // DHT
		int L = D.be16() - 2;
		while (L > 0) {
			const int q = D.byte(), tc = q >> 4, th = q & 15;
			if (tc > 1 || th > 3) return D.fail("JPEG: bad DHT");
			int counts[16], n = 0;
			for (int i = 0; i < 16; i++) { counts[i] = D.byte(); n += counts[i]; }
			if (n > 256) return D.fail("JPEG: bad DHT");
			Huff& h = tc ? D.ac[th] : D.dc[th];
			if (!build_huff(h, counts)) return D.fail("JPEG: bad Huffman code lengths");
			for (int i = 0; i < n; i++) h.values[i] = (uint8_t)D.byte();
			L -= 17 + n;
		}
		return L == 0 ? true : D.fail("JPEG: bad DHT length");
	}
	if ((m >= 0xe0 && m <= 0xef) || m == 0xfe) { // APPn / COM
		int L = D.be16();
		if (L < 2) return D.fail("JPEG: bad segment length");
		L -= 2;
		if (m == 0xe0 && L >= 5) {
			static const char tag[5] = { 'J', 'F', 'I', 'F', 0 };
			bool ok = true;
			for (int i = 0; i < 5; i++) if (D.byte() != (uint8_t)tag[i]) ok = false;
			L -= 5;
			if (ok) D.jfif = true;
		}
		else if (m == 0xee && L >= 12) {
			static const char tag[6] = { 'A', 'd', 'o', 'b', 'e', 0 };
			bool ok = true;
			for (int i = 0; i < 6; i++) if (D.byte() != (uint8_t)tag[i]) ok = false;
			L -= 6;
			if (ok) {
				D.byte(); D.be16(); D.be16();
				D.adobe_transform = D.byte();
				L -= 6;
			}
		}
		if (L > D.end - D.p) return D.fail("JPEG: truncated segment");
		D.p += L;
		return true;
	}
	return D.fail("JPEG: unknown marker");
}

int next_marker(Decoder& D)
{
	if (D.marker) { const int m = D.marker; D.marker = 0; return m; }
	int x = D.byte();
	if (x != 0xff) return 0;
	while (x == 0xff) x = D.byte();
	return x;
}

bool read_frame_header(Decoder& D)
{
	const int Lf = D.be16();
	if (Lf < 11) return D.fail("JPEG: bad SOF length");
	if (D.byte() != 8) return D.fail("JPEG: only 8-bit samples are supported");
	D.height = D.be16();
	D.width = D.be16();
	if (D.height == 0 || D.width == 0) return D.fail("JPEG: zero image size");
	if (D.width > (1 << 24) || D.height > (1 << 24)) return D.fail("JPEG: image too large");
	D.ncomp = D.byte();
	if (D.ncomp != 1 && D.ncomp != 3 && D.ncomp != 4) return D.fail("JPEG: bad component count");
	if (Lf != 8 + 3 * D.ncomp) return D.fail("JPEG: bad SOF length");
	D.rgb_ids = 0;
	for (int i = 0; i < D.ncomp; i++) {
		Component& c = D.comp[i];
		c.id = D.byte();
		if (D.ncomp == 3 && c.id == "RGB"[i]) ++D.rgb_ids;
		const int q = D.byte();
		c.h = q >> 4; c.v = q & 15;
		if (c.h < 1 || c.h > 4 || c.v < 1 || c.v > 4) return D.fail("JPEG: bad sampling factor");
		c.tq = D.byte();
		if (c.tq > 3) return D.fail("JPEG: bad quantisation table index");
		if (c.h > D.h_max) D.h_max = c.h;
		if (c.v > D.v_max) D.v_max = c.v;
	}
	if ((uint64_t)D.width * D.height * D.ncomp > (1ull << 31)) return D.fail("JPEG: image too large");
	for (int i = 0; i < D.ncomp; i++)
		if (D.h_max % D.comp[i].h || D.v_max % D.comp[i].v) return D.fail("JPEG: fractional sampling ratios are not supported");
	D.mcu_x = (D.width + D.h_max * 8 - 1) / (D.h_max * 8);
	D.mcu_y = (D.height + D.v_max * 8 - 1) / (D.v_max * 8);
	for (int i = 0; i < D.ncomp; i++) {
		Component& c = D.comp[i];
		c.x = (D.width * c.h + D.h_max - 1) / D.h_max;
		c.y = (D.height * c.v + D.v_max - 1) / D.v_max;
		c.w2 = D.mcu_x * c.h * 8;
		c.h2 = D.mcu_y * c.v * 8;
		c.plane.assign((size_t)c.w2 * c.h2, 0);
		if (D.progressive) {
			c.blocks_w = c.w2 / 8;
			c.coeff.assign((size_t)c.w2 * c.h2, 0);
		}
	}
	return true;
}

bool read_scan_header(Decoder& D)
{
	const int Ls = D.be16();
	D.scan_n = D.byte();
	if (D.scan_n < 1 || D.scan_n > 4 || D.scan_n > D.ncomp) return D.fail("JPEG: bad SOS component count");
	if (Ls != 6 + 2 * D.scan_n) return D.fail("JPEG: bad SOS length");
	for (int i = 0; i < D.scan_n; i++) {
		const int id = D.byte(), q = D.byte();
		int which = 0;
		while (which < D.ncomp && D.comp[which].id != id) which++;
		if (which == D.ncomp) return D.fail("JPEG: SOS names an unknown component");
		D.comp[which].hd = q >> 4; D.comp[which].ha = q & 15;
		if (D.comp[which].hd > 3 || D.comp[which].ha > 3) return D.fail("JPEG: bad Huffman table index");
		D.order[i] = which;
	}
	D.ss = D.byte(); D.se = D.byte();
	const int a = D.byte();
	D.ah = a >> 4; D.al = a & 15;
	if (D.progressive) {
		if (D.ss > 63 || D.se > 63 || D.ss > D.se || D.ah > 13 || D.al > 13) return D.fail("JPEG: bad SOS");
	}
	else {
		if (D.ss != 0 || D.ah != 0 || D.al != 0) return D.fail("JPEG: bad SOS");
		D.se = 63;
	}
	return true;
}

// Junk after a scan: look for the next thing that is a marker (stb_image.h:3388-3409).
int skip_to_marker(Decoder& D)
{
	while (!D.at_end()) {
		int x = D.byte();
		while (x == 0xff) {
			if (D.at_end()) return 0;
			x = D.byte();
			if (x != 0x00 && x != 0xff) return x;
		}
	}
	return 0;
}

bool decode_frames(Decoder& D)
{
	if (next_marker(D) != 0xd8) return D.fail("JPEG: no SOI marker");
	int m = next_marker(D);
	while (!(m == 0xc0 || m == 0xc1 || m == 0xc2)) {
		if (m == 0) return D.fail("JPEG: expected a marker");
		if (!read_tables(D, m)) return false;
		m = next_marker(D);
		while (m == 0) {
			if (D.at_end()) return D.fail("JPEG: no SOF marker");
			m = next_marker(D);
		}
	}
	D.progressive = m == 0xc2;
	if (!read_frame_header(D)) return false;
	m = next_marker(D);
	while (m != 0xd9) {
		if (m == 0xda) {
			if (!read_scan_header(D)) return false;
			if (!decode_scan(D)) return false;
			if (D.marker == 0) D.marker = skip_to_marker(D);
			m = next_marker(D);
			if (m >= 0xd0 && m <= 0xd7) m = next_marker(D);
		}
		else if (m == 0xdc) { // DNL
			const int Ld = D.be16(), NL = D.be16();
			if (Ld != 4 || NL != D.height) return D.fail("JPEG: bad DNL");
			m = next_marker(D);
		}
		else {
			if (m == 0 || !read_tables(D, m)) { D.err.clear(); break; } // stb keeps what it has (stb_image.h:3446)
			m = next_marker(D);
		}
		if (m == 0 && D.at_end()) break;
	}
	if (D.progressive) // dequantise in 16 bits, then the inverse DCT (stb_image.h:3072-3096)
		for (int n = 0; n < D.ncomp; n++) {
			Component& c = D.comp[n];
			const uint16_t* q = D.quant[c.tq];
			const int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3;
			for (int j = 0; j < bh; j++)
				for (int i = 0; i < bw; i++) {
					int16_t* data = c.coeff.data() + 64 * ((size_t)i + (size_t)j * c.blocks_w);
					for (int k = 0; k < 64; k++) data[k] = (int16_t)(data[k] * q[k]);
					idct_block(c.plane.data() + (size_t)c.w2 * j * 8 + i * 8, c.w2, data);
				}
		}
	return true;
}

// ---- upsampling of one output row from the two nearest rows of a subsampled plane (stb_image.h:3464-3657) ----
const uint8_t* up_v2(uint8_t* out, const uint8_t* nearr, const uint8_t* farr, int w)
{
	for (int i = 0; i < w; i++) out[i] = (uint8_t)((3 * nearr[i] + farr[i] + 2) >> 2);
	return out;
}
const uint8_t* up_h2(uint8_t* out, const uint8_t* in, int w)
{
	if (w == 1) { out[0] = out[1] = in[0]; return out; }
	out[0] = in[0];
	out[1] = (uint8_t)((in[0] * 3 + in[1] + 2) >> 2);
	int i;
	for (i = 1; i < w - 1; i++) {
		const int n = 3 * in[i] + 2;
		out[2 * i] = (uint8_t)((n + in[i - 1]) >> 2);
		out[2 * i + 1] = (uint8_t)((n + in[i + 1]) >> 2);
	}
	out[2 * i] = (uint8_t)((in[w - 2] * 3 + in[w - 1] + 2) >> 2);
	out[2 * i + 1] = in[w - 1];
	return out;
}
const uint8_t* up_hv2(uint8_t* out, const uint8_t* nearr, const uint8_t* farr, int w)
{
	if (w == 1) { out[0] = out[1] = (uint8_t)((3 * nearr[0] + farr[0] + 2) >> 2); return out; }
	int t1 = 3 * nearr[0] + farr[0];
	out[0] = (uint8_t)((t1 + 2) >> 2);
	for (int i = 1; i < w; i++) {
		const int t0 = t1;
		t1 = 3 * nearr[i] + farr[i];
		out[2 * i - 1] = (uint8_t)((3 * t0 + t1 + 8) >> 4);
		out[2 * i] = (uint8_t)((3 * t1 + t0 + 8) >> 4);
	}
	out[2 * w - 1] = (uint8_t)((t1 + 2) >> 2);
	return out;
}
const uint8_t* up_nearest(uint8_t* out, const uint8_t* in, int w, int hs)
{
	for (int i = 0; i < w; i++) for (int j = 0; j < hs; j++) out[i * hs + j] = in[i];
	return out;
}

inline uint8_t mul8(uint8_t x, uint8_t y) { const unsigned t = (unsigned)x * y + 128; return (uint8_t)((t + (t >> 8)) >> 8); } // stb_image.h:3858

void ycc_row(uint8_t* out, const uint8_t* y, const uint8_t* cb, const uint8_t* cr, int n, int step)
{
	for (int i = 0; i < n; i++, out += step) {
		const int yf = (y[i] << 20) + (1 << 19);
		const int r8 = cr[i] - 128, b8 = cb[i] - 128;
		int r = yf + r8 * 1470208;                                                        // 1.40200 -> 5743 << 8
		int g = yf + r8 * -748800 + (int)((unsigned)(b8 * -360960) & 0xffff0000u);        // 0.71414 -> 2925 << 8, 0.34414 -> 1410 << 8
		int b = yf + b8 * 1858048;                                                        // 1.77200 -> 7258 << 8
		out[0] = clamp255(r >> 20); out[1] = clamp255(g >> 20); out[2] = clamp255(b >> 20);
	}
}

} // namespace

bool looks_like_jpeg(const uint8_t* data, size_t len) { return len >= 3 && data[0] == 0xff && data[1] == 0xd8 && data[2] == 0xff; }

bool decode_jpeg(const uint8_t* data, size_t len, HostTexture& out, std::string& err)
{
	Decoder D;
	D.p = data; D.end = data + len;
	memset(D.quant, 0, sizeof(D.quant));
	if (!decode_frames(D)) { err = D.err.empty() ? "JPEG: corrupt stream" : D.err; return false; }
	const int W = D.width, H = D.height;
	const int n = D.ncomp >= 3 ? 3 : 1; // req_comp = 0: 3 channels for colour, 1 for grey (stb_image.h:3877, 4018)
	const bool is_rgb = D.ncomp == 3 && (D.rgb_ids == 3 || (D.adobe_transform == 0 && !D.jfif));
	out.width = W; out.height = H; out.comps = n;
	out.pixels.assign((size_t)W * H * n, 0);
	struct Up { int hs, vs, ystep, w_lores, ypos; const uint8_t* line0; const uint8_t* line1; std::vector<uint8_t> buf; } up[4];
	for (int k = 0; k < D.ncomp; k++) {
		Up& u = up[k];
		u.hs = D.h_max / D.comp[k].h; u.vs = D.v_max / D.comp[k].v;
		u.ystep = u.vs >> 1;
		u.w_lores = (W + u.hs - 1) / u.hs;
		u.ypos = 0;
		u.line0 = u.line1 = D.comp[k].plane.data();
		u.buf.assign((size_t)W + 3 + 8, 0);
	}
	const uint8_t* row[4] = { nullptr, nullptr, nullptr, nullptr };
	for (int j = 0; j < H; j++) {
		uint8_t* o = out.pixels.data() + (size_t)n * W * j;
		for (int k = 0; k < D.ncomp; k++) {
			Up& u = up[k];
			const bool bottom = u.ystep >= (u.vs >> 1);
			const uint8_t* nearr = bottom ? u.line1 : u.line0;
			const uint8_t* farr = bottom ? u.line0 : u.line1;
			if (u.hs == 1 && u.vs == 1) row[k] = nearr;
			else if (u.hs == 1 && u.vs == 2) row[k] = up_v2(u.buf.data(), nearr, farr, u.w_lores);
			else if (u.hs == 2 && u.vs == 1) row[k] = up_h2(u.buf.data(), nearr, u.w_lores);
			else if (u.hs == 2 && u.vs == 2) row[k] = up_hv2(u.buf.data(), nearr, farr, u.w_lores);
			else row[k] = up_nearest(u.buf.data(), nearr, u.w_lores, u.hs);
			if (++u.ystep >= u.vs) {
				u.ystep = 0;
				u.line0 = u.line1;
				if (++u.ypos < D.comp[k].y) u.line1 += D.comp[k].w2;
			}
		}
		if (D.ncomp == 1) memcpy(o, row[0], (size_t)W);
		else if (D.ncomp == 3) {
			if (is_rgb) for (int i = 0; i < W; i++, o += 3) { o[0] = row[0][i]; o[1] = row[1][i]; o[2] = row[2][i]; }
			else ycc_row(o, row[0], row[1], row[2], W, 3);
		}
		else if (D.adobe_transform == 0) // CMYK
			for (int i = 0; i < W; i++, o += 3) { const uint8_t m = row[3][i]; o[0] = mul8(row[0][i], m); o[1] = mul8(row[1][i], m); o[2] = mul8(row[2][i], m); }
		else if (D.adobe_transform == 2) { // YCCK
			ycc_row(o, row[0], row[1], row[2], W, 3);
			for (int i = 0; i < W; i++, o += 3) { const uint8_t m = row[3][i]; o[0] = mul8(255 - o[0], m); o[1] = mul8(255 - o[1], m); o[2] = mul8(255 - o[2], m); }
		}
		else ycc_row(o, row[0], row[1], row[2], W, 3); // YCbCr + an ignored fourth channel
	}
	return true;
}

// What the reference hands to stbi_load_from_memory (Texture.cu:21-30): stb tries PNG first, JPEG later (stb_image.h:1141-1170);
// the other containers stb knows (BMP, GIF, PSD, PIC, PNM, HDR, TGA) do not occur in glTF (the specification allows PNG and JPEG).
bool decode_image(const uint8_t* data, size_t len, HostTexture& out, std::string& err)
{
	if (looks_like_jpeg(data, len)) return decode_jpeg(data, len, out, err);
	if (!decode_png(data, len, out, err)) {
		if (err.find("not a PNG") != std::string::npos) err = "not a PNG or JPEG image";
		return false;
	}
	return true;
}

} // namespace drt
