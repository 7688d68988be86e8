// drt_png.cpp -- PNG decoder (zlib inflate + PNG unfiltering) for scene textures.
//
// The reference decodes embedded images with stb_image 2.29 and req_comp = 0
// (DustRayTracer/src/Core/Scene/Texture.cu:21-30), i.e. it keeps the file's own channel count, which
// later decides whether a texture is alpha-cut-out capable (AnyHit.cuh:17). This decoder reproduces the
// channel-count and 8-bit conversion conventions of that call:
//   gray -> 1, gray+alpha -> 2, RGB -> 3, RGBA -> 4, palette -> 3 (4 with tRNS);
//   tRNS on gray/RGB adds an alpha channel (0 for the key colour, 255 otherwise);
//   1/2/4-bit gray is scaled to 0..255; 16-bit samples keep their high byte.
#include "drt_host.h"

#include <zlib.h>
#include <cstring>

namespace drt {

namespace {

uint32_t be32(const uint8_t* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

int paeth(int a, int b, int c)
{
	int p = a + b - c;
	int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
	if (pa <= pb && pa <= pc) return a;
	if (pb <= pc) return b;
	return c;
}

// Unfilters one (sub)image of w x h pixels, `bpp_bits` bits per pixel, from `src` into `dst`
// (dst row stride = ceil(w*bpp_bits/8), no filter bytes). Returns bytes consumed or 0 on error.
size_t unfilter(const uint8_t* src, size_t src_len, uint8_t* dst, uint32_t w, uint32_t h, uint32_t bpp_bits)
{
	const size_t stride = ((size_t)w * bpp_bits + 7) / 8;
	const size_t bpp = (bpp_bits + 7) / 8;
	if (src_len < (stride + 1) * h) return 0;
	const uint8_t* prev = nullptr;
	for (uint32_t y = 0; y < h; y++) {
		const uint8_t ft = *src++;
		uint8_t* row = dst + (size_t)y * stride;
		for (size_t i = 0; i < stride; i++) {
			int a = i >= bpp ? row[i - bpp] : 0;
			int b = prev ? prev[i] : 0;
			int c = (prev && i >= bpp) ? prev[i - bpp] : 0;
			int x = src[i];
			switch (ft) {
			case 0: break;
			case 1: x += a; break;
			case 2: x += b; break;
			case 3: x += (a + b) >> 1; break;
			case 4: x += paeth(a, b, c); break;
			default: return 0;
			}
			row[i] = (uint8_t)x;
		}
		src += stride;
		prev = row;
	}
	return (stride + 1) * h;
}

} // namespace

bool decode_png(const uint8_t* data, size_t len, HostTexture& out, std::string& err)
{
	static const uint8_t sig[8] = { 0x89, 'P', 'N', 'G', '\r', '\n', 0x1a, '\n' };
	if (len < 8 || memcmp(data, sig, 8) != 0) { err = "not a PNG image (only PNG textures are decoded)"; return false; }
	uint32_t w = 0, h = 0;
	int depth = 0, color = 0, interlace = 0;
	std::vector<uint8_t> idat;
	uint8_t palette[256][4];
	int pal_len = 0;
	bool has_trns = false;
	uint16_t trns_key[3] = { 0, 0, 0 };
	for (int i = 0; i < 256; i++) { palette[i][0] = palette[i][1] = palette[i][2] = 0; palette[i][3] = 255; }
	size_t p = 8;
	bool seen_ihdr = false, seen_iend = false;
	while (p + 12 <= len && !seen_iend) {
		uint32_t clen = be32(data + p);
		const uint8_t* tag = data + p + 4;
		const uint8_t* body = data + p + 8;
		if (p + 12 + (size_t)clen > len) { err = "truncated PNG chunk"; return false; }
		if (!memcmp(tag, "IHDR", 4)) {
			if (clen < 13) { err = "bad IHDR"; return false; }
			w = be32(body); h = be32(body + 4);
			depth = body[8]; color = body[9]; interlace = body[12];
			if (body[10] != 0 || body[11] != 0 || interlace > 1) { err = "unsupported PNG method"; return false; }
			seen_ihdr = true;
		}
		else if (!memcmp(tag, "PLTE", 4)) {
			pal_len = (int)(clen / 3);
			if (pal_len > 256) pal_len = 256;
			for (int i = 0; i < pal_len; i++) { palette[i][0] = body[3 * i]; palette[i][1] = body[3 * i + 1]; palette[i][2] = body[3 * i + 2]; }
		}
		else if (!memcmp(tag, "tRNS", 4)) {
			has_trns = true;
			if (color == 3) { for (uint32_t i = 0; i < clen && i < 256; i++) palette[i][3] = body[i]; }
			else if (color == 0 && clen >= 2) trns_key[0] = (uint16_t)((body[0] << 8) | body[1]);
			else if (color == 2 && clen >= 6) for (int k = 0; k < 3; k++) trns_key[k] = (uint16_t)((body[2 * k] << 8) | body[2 * k + 1]);
			else has_trns = false;
		}
		else if (!memcmp(tag, "IDAT", 4)) idat.insert(idat.end(), body, body + clen);
		else if (!memcmp(tag, "IEND", 4)) seen_iend = true;
		p += 12 + (size_t)clen;
	}
	if (!seen_ihdr || w == 0 || h == 0 || idat.empty()) { err = "PNG without IHDR/IDAT"; return false; }
	int src_ch;
	switch (color) {
	case 0: src_ch = 1; break;
	case 2: src_ch = 3; break;
	case 3: src_ch = 1; break;
	case 4: src_ch = 2; break;
	case 6: src_ch = 4; break;
	default: err = "bad PNG colour type"; return false;
	}
	if (!(depth == 1 || depth == 2 || depth == 4 || depth == 8 || depth == 16) || (color == 3 && depth == 16)
		|| ((color == 2 || color == 4 || color == 6) && depth < 8)) { err = "bad PNG bit depth"; return false; }
	if ((uint64_t)w * h > (1ull << 28)) { err = "PNG too large"; return false; }
	const uint32_t bpp_bits = (uint32_t)(src_ch * depth);
	const size_t stride = ((size_t)w * bpp_bits + 7) / 8;

	// inflate
	size_t raw_cap = (stride + 1) * (size_t)h + (interlace ? 8 * (size_t)h + 64 : 0);
	std::vector<uint8_t> raw(raw_cap);
	z_stream zs;
	memset(&zs, 0, sizeof(zs));
	if (inflateInit(&zs) != Z_OK) { err = "zlib init failed"; return false; }
	zs.next_in = idat.data(); zs.avail_in = (uInt)idat.size();
	zs.next_out = raw.data(); zs.avail_out = (uInt)raw.size();
	int zr = inflate(&zs, Z_FINISH);
	size_t raw_len = zs.total_out;
	inflateEnd(&zs);
	if (zr != Z_STREAM_END && zr != Z_OK && zr != Z_BUF_ERROR) { err = "PNG inflate failed"; return false; }

	// unfilter (+ de-interlace) into packed rows of the source format
	std::vector<uint8_t> img(stride * h);
	if (!interlace) {
		if (!unfilter(raw.data(), raw_len, img.data(), w, h, bpp_bits)) { err = "PNG data too short / bad filter"; return false; }
	}
	else {
		static const int xs[7] = { 0, 4, 0, 2, 0, 1, 0 }, ys[7] = { 0, 0, 4, 0, 2, 0, 1 };
		static const int dx[7] = { 8, 8, 4, 4, 2, 2, 1 }, dy[7] = { 8, 8, 8, 4, 4, 2, 2 };
		size_t off = 0;
		for (int pass = 0; pass < 7; pass++) {
			uint32_t pw = (w - xs[pass] + dx[pass] - 1) / dx[pass], ph = (h - ys[pass] + dy[pass] - 1) / dy[pass];
			if ((int)w <= xs[pass] || (int)h <= ys[pass] || pw == 0 || ph == 0) continue;
			size_t pstride = ((size_t)pw * bpp_bits + 7) / 8;
			std::vector<uint8_t> sub(pstride * ph);
			size_t used = unfilter(raw.data() + off, raw_len - off, sub.data(), pw, ph, bpp_bits);
			if (!used) { err = "PNG interlace pass too short"; return false; }
			off += used;
			for (uint32_t y = 0; y < ph; y++)
				for (uint32_t x = 0; x < pw; x++) {
					uint32_t X = xs[pass] + x * dx[pass], Y = ys[pass] + y * dy[pass];
					if (bpp_bits >= 8) memcpy(&img[Y * stride + (size_t)X * (bpp_bits / 8)], &sub[y * pstride + (size_t)x * (bpp_bits / 8)], bpp_bits / 8);
					else {
						uint32_t sbit = x * bpp_bits, dbit = X * bpp_bits;
						uint8_t v = (uint8_t)((sub[y * pstride + sbit / 8] >> (8 - bpp_bits - (sbit & 7))) & ((1u << bpp_bits) - 1u));
						uint8_t& d = img[Y * stride + dbit / 8];
						int sh = 8 - (int)bpp_bits - (int)(dbit & 7);
						d = (uint8_t)((d & ~(((1u << bpp_bits) - 1u) << sh)) | (v << sh));
					}
				}
		}
	}

	// expand to 8-bit interleaved with stb's channel conventions
	int out_ch;
	if (color == 3) out_ch = has_trns ? 4 : 3;
	else out_ch = src_ch + ((has_trns && (color == 0 || color == 2)) ? 1 : 0);
	out.width = (int)w; out.height = (int)h; out.comps = out_ch;
	out.pixels.assign((size_t)w * h * out_ch, 0);
	static const int gray_scale[9] = { 0, 0xff, 0x55, 0, 0x11, 0, 0, 0, 0x01 };
	for (uint32_t y = 0; y < h; y++) {
		const uint8_t* row = img.data() + (size_t)y * stride;
		uint8_t* o = out.pixels.data() + (size_t)y * w * out_ch;
		for (uint32_t x = 0; x < w; x++) {
			uint16_t s[4] = { 0, 0, 0, 0 };
			if (depth == 16) for (int c = 0; c < src_ch; c++) s[c] = (uint16_t)((row[((size_t)x * src_ch + c) * 2] << 8) | row[((size_t)x * src_ch + c) * 2 + 1]);
			else if (depth == 8) for (int c = 0; c < src_ch; c++) s[c] = row[(size_t)x * src_ch + c];
			else { uint32_t bit = x * depth; s[0] = (uint16_t)((row[bit / 8] >> (8 - depth - (bit & 7))) & ((1u << depth) - 1u)); }
			if (color == 3) {
				const uint8_t* pe = palette[s[0] & 0xff];
				o[0] = pe[0]; o[1] = pe[1]; o[2] = pe[2];
				if (out_ch == 4) o[3] = pe[3];
			}
			else {
				bool key = has_trns && (color == 0 ? s[0] == trns_key[0] : (color == 2 && s[0] == trns_key[0] && s[1] == trns_key[1] && s[2] == trns_key[2]));
				for (int c = 0; c < src_ch; c++) {
					uint8_t v;
					if (depth == 16) v = (uint8_t)(s[c] >> 8);
					else if (depth == 8) v = (uint8_t)s[c];
					else v = (uint8_t)(s[c] * gray_scale[depth]);
					o[c] = v;
				}
				if (out_ch == src_ch + 1) o[src_ch] = key ? 0 : 255;
			}
			o += out_ch;
		}
	}
	return true;
}

} // namespace drt
