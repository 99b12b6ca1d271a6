/*
 * TEST INFRASTRUCTURE ONLY -- see snappy_oracle.h.
 *
 * Index-based restatement of the reference host codec.  Written from the
 * behaviour documented in SURVEY.md section 8(a) and checked against the
 * reference itself (oracle/_ref) -- not a copy of its source.
 */
#include "snappy_oracle.h"

#include <string.h>

/* ------------------------------------------------------------------ util */

static uint32_t le32(const uint8_t *p)
{
	return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
}

static void put_le32(uint8_t *p, uint32_t v)
{
	p[0] = (uint8_t)v;
	p[1] = (uint8_t)(v >> 8);
	p[2] = (uint8_t)(v >> 16);
	p[3] = (uint8_t)(v >> 24);
}

/* snappy_compress.c:69-98 -- minimal little-endian base-128 varint */
static size_t put_varint32(uint8_t *p, uint32_t v)
{
	size_t k = 0;
	while (v >= 0x80) {
		p[k++] = (uint8_t)(v | 0x80);
		v >>= 7;
	}
	p[k++] = (uint8_t)v;
	return k;
}

/* snappy_decompress.c:23-37 -- at most five bytes; 0 = malformed */
static size_t get_varint32(const uint8_t *p, size_t n, uint32_t *v)
{
	uint32_t acc = 0;
	for (size_t k = 0; k < 5 && k < n; k++) {
		acc |= (uint32_t)(p[k] & 0x7F) << (7 * k);
		if (!(p[k] & 0x80)) {
			*v = acc;
			return k + 1;
		}
	}
	return 0;
}

size_t oracle_max_compressed_length(size_t n)
{
	return n ? 32 + n + n / 6 : 0; /* snappy_compress.c:55-60 */
}

size_t oracle_file_bound(size_t n, uint32_t bs)
{
	size_t nblocks = bs ? (n + bs - 1) / bs : 0;
	return 16 + nblocks * (4 + 8) + n + n / 6 + 32;
}

size_t oracle_read_header(const uint8_t *in, size_t n, uint32_t *dlen, uint32_t *bs)
{
	size_t a = get_varint32(in, n, dlen);
	if (!a)
		return 0;
	size_t b = get_varint32(in + a, n - a, bs);
	if (!b)
		return 0;
	return a + b;
}

/* -------------------------------------------------------------- compress */

#define TABLE_MAX 16384u /* snappy_compress.c:16-17 */

/* snappy_compress.c:202-225 */
static size_t put_literal(uint8_t *o, const uint8_t *src, uint32_t len)
{
	uint32_t n = len - 1;
	size_t k = 0;
	if (n < 60) {
		o[k++] = (uint8_t)(n << 2);
	} else {
		uint32_t extra = 0;
		for (uint32_t t = n; t; t >>= 8)
			o[1 + extra++] = (uint8_t)t;
		o[0] = (uint8_t)((59 + extra) << 2);
		k = 1 + extra;
	}
	memcpy(o + k, src, len);
	return k + len;
}

/* snappy_compress.c:234-245 */
static size_t put_copy_upto64(uint8_t *o, uint32_t off, uint32_t len)
{
	if (len < 12 && off < 2048) {
		o[0] = (uint8_t)(1 + ((len - 4) << 2) + ((off >> 8) << 5));
		o[1] = (uint8_t)off;
		return 2;
	}
	o[0] = (uint8_t)(2 + ((len - 1) << 2));
	o[1] = (uint8_t)off;
	o[2] = (uint8_t)(off >> 8); /* 16-bit truncation is the reference's, :241-243 */
	return 3;
}

/* snappy_compress.c:254-272 */
static size_t put_copy(uint8_t *o, uint32_t off, uint32_t len)
{
	size_t k = 0;
	while (len >= 68) {
		k += put_copy_upto64(o + k, off, 64);
		len -= 64;
	}
	if (len > 64) {
		k += put_copy_upto64(o + k, off, 60);
		len -= 60;
	}
	return k + put_copy_upto64(o + k, off, len);
}

size_t oracle_compress_block(const uint8_t *in, uint32_t n, uint8_t *out)
{
	uint16_t table[TABLE_MAX];
	/* snappy_compress.c:139-146: power of two in [256, 16384], >= n if possible */
	uint32_t tsize = 256;
	while (tsize < TABLE_MAX && tsize < n)
		tsize <<= 1;
	memset(table, 0, tsize * sizeof(table[0]));
	int lg = 0;
	while ((1u << (lg + 1)) <= tsize)
		lg++;
	const int shift = 32 - lg; /* :288 */
#define H(pos) ((le32(in + (pos)) * 0x1e35a7bdu) >> shift) /* :161-166 */

	uint8_t *o = out + 4;
	uint32_t emit_from = 0; /* first byte not yet covered by an element */

	if (n >= 15) { /* :299-301 */
		const uint32_t limit = n - 15;
		uint32_t ip = 1;
		uint32_t hnext = H(ip);
		for (;;) {
			/* step 1 (:333-348): probe with the growing-stride heuristic */
			uint32_t skip = 32, cand, nxt = ip;
			int out_of_input = 0;
			do {
				ip = nxt;
				uint32_t h = hnext;
				nxt = ip + (skip++ >> 5);
				if (nxt > limit) {
					out_of_input = 1;
					break;
				}
				hnext = H(nxt);
				cand = table[h];
				table[h] = (uint16_t)ip;
			} while (le32(in + ip) != le32(in + cand));
			if (out_of_input)
				break;

			/* step 2 (:355) */
			o += put_literal(o, in + emit_from, ip - emit_from);

			/* step 3 (:370-398) */
			int done = 0;
			uint32_t cbytes;
			do {
				uint32_t start = ip;
				uint32_t m = 4; /* :176-193, bounded by the end of the block */
				while (ip + m < n && in[cand + m] == in[ip + m])
					m++;
				ip += m;
				o += put_copy(o, start - cand, m);
				emit_from = ip;
				if (ip >= limit) {
					done = 1;
					break;
				}
				table[H(ip - 1)] = (uint16_t)(ip - 1);
				uint32_t h = H(ip);
				cand = table[h];
				cbytes = le32(in + cand);
				table[h] = (uint16_t)ip;
			} while (le32(in + ip) == cbytes);
			if (done)
				break;
			hnext = H(ip + 1); /* :400-401 */
			ip++;
		}
	}
#undef H
	if (emit_from < n) /* :405-410 */
		o += put_literal(o, in + emit_from, n - emit_from);
	put_le32(out, (uint32_t)(o - (out + 4))); /* :412 */
	return (size_t)(o - out);
}

int oracle_compress(const uint8_t *in, size_t n, uint32_t bs, uint8_t *out, size_t cap, size_t *out_len)
{
	if (bs == 0 || cap < oracle_file_bound(n, bs))
		return ORACLE_BUFFER_TOO_SMALL;
	size_t o = put_varint32(out, (uint32_t)n); /* :461-465 */
	o += put_varint32(out + o, bs);
	for (size_t pos = 0; pos < n;) { /* :467-479 */
		uint32_t take = (uint32_t)((n - pos) < bs ? (n - pos) : bs);
		o += oracle_compress_block(in + pos, take, out + o);
		pos += take;
	}
	*out_len = o;
	return ORACLE_OK;
}

/* ------------------------------------------------------------ decompress */

int oracle_decompress(const uint8_t *in, size_t n, uint8_t *out, size_t cap, size_t *out_len)
{
	uint32_t dlen, bs;
	size_t ip = get_varint32(in, n, &dlen); /* setup_decompression :194-198 */
	if (!ip)
		return ORACLE_INVALID_INPUT;
	if (dlen > cap) /* :201-204 */
		return ORACLE_BUFFER_TOO_SMALL;
	size_t k = get_varint32(in + ip, n - ip, &bs); /* :221-225; value unused */
	if (!k)
		return ORACLE_INVALID_INPUT;
	ip += k;
	*out_len = dlen;
	size_t op = 0;

	while (ip < n) { /* :227 */
		uint32_t csize = le32(in + ip);
		ip += 4;
		size_t bend = ip + csize;
		while (ip != bend) { /* :232 */
			uint8_t tag = in[ip++];
			uint16_t length; /* 16 bits, as in the reference (:233) */
			uint32_t offset = 0;
			uint32_t six = tag >> 2;
			int is_copy = 1;
			switch (tag & 3) {
			case 0:
				is_copy = 0;
				length = (uint16_t)(six + 1);
				if (length > 60) {
					uint32_t nb = length - 60, v = 0; /* :64-74 */
					if (ip + nb < n) {
						for (uint32_t i = 0; i < nb; i++)
							v |= (uint32_t)in[ip++] << (8 * i);
					}
					length = (uint16_t)(v + 1);
				}
				break;
			case 1:
				length = (uint16_t)(((tag >> 2) & 7) + 4);
				if (ip < n) /* :83-88 */
					offset = in[ip++] | ((uint32_t)(tag >> 5) << 8);
				break;
			case 2:
				length = (uint16_t)(six + 1);
				if (ip + 2 <= n) { /* :97-109 */
					offset = in[ip] | ((uint32_t)in[ip + 1] << 8);
					ip += 2;
				}
				break;
			default:
				length = (uint16_t)(six + 1);
				if (ip + 4 <= n) { /* :118-133 */
					offset = le32(in + ip);
					ip += 4;
				}
				break;
			}
			if (!is_copy) { /* :142-154 */
				uint32_t l = length;
				while (l && ip < n && op < dlen) {
					out[op++] = in[ip++];
					l--;
				}
			} else { /* :164-184 */
				if (offset > op)
					return ORACLE_INVALID_INPUT;
				uint32_t l = length;
				size_t s = op - offset;
				while (l && op < dlen) {
					out[op++] = out[s++];
					l--;
				}
			}
			if (ip > bend) /* the reference would run away here; stop instead */
				return ORACLE_INVALID_INPUT;
		}
	}
	return ORACLE_OK;
}

int oracle_decompress_strict(const uint8_t *in, size_t n, uint8_t *out, size_t cap, size_t *out_len)
{
	uint32_t dlen, bs;
	size_t ip = oracle_read_header(in, n, &dlen, &bs);
	if (!ip)
		return ORACLE_INVALID_INPUT;
	if (dlen > cap)
		return ORACLE_BUFFER_TOO_SMALL;
	if (bs == 0 && dlen != 0)
		return ORACLE_INVALID_INPUT;
	*out_len = dlen;
	size_t done = 0;
	while (done < dlen) {
		if (ip + 4 > n)
			return ORACLE_INVALID_INPUT;
		uint32_t csize = le32(in + ip);
		ip += 4;
		if (csize > n - ip)
			return ORACLE_INVALID_INPUT;
		const uint8_t *c = in + ip;
		uint8_t *b = out + done;
		size_t want = dlen - done < bs ? dlen - done : bs;
		size_t cp = 0, bp = 0;
		while (cp < csize) {
			uint8_t tag = c[cp++];
			uint32_t len, off;
			switch (tag & 3) {
			case 0: {
				len = (tag >> 2) + 1;
				if (len > 60) {
					uint32_t nb = len - 60;
					if (cp + nb > csize)
						return ORACLE_INVALID_INPUT;
					len = 0;
					for (uint32_t i = 0; i < nb; i++)
						len |= (uint32_t)c[cp++] << (8 * i);
					if (len == 0xFFFFFFFFu)
						return ORACLE_INVALID_INPUT;
					len += 1;
				}
				if (len > csize - cp || len > want - bp)
					return ORACLE_INVALID_INPUT;
				memcpy(b + bp, c + cp, len);
				cp += len;
				bp += len;
				continue;
			}
			case 1:
				if (cp + 1 > csize)
					return ORACLE_INVALID_INPUT;
				len = ((tag >> 2) & 7) + 4;
				off = c[cp] | ((uint32_t)(tag >> 5) << 8);
				cp += 1;
				break;
			case 2:
				if (cp + 2 > csize)
					return ORACLE_INVALID_INPUT;
				len = (tag >> 2) + 1;
				off = c[cp] | ((uint32_t)c[cp + 1] << 8);
				cp += 2;
				break;
			default:
				if (cp + 4 > csize)
					return ORACLE_INVALID_INPUT;
				len = (tag >> 2) + 1;
				off = le32(c + cp);
				cp += 4;
				break;
			}
			if (off == 0 || off > bp || len > want - bp)
				return ORACLE_INVALID_INPUT;
			for (uint32_t i = 0; i < len; i++, bp++)
				b[bp] = b[bp - off];
		}
		if (bp != want)
			return ORACLE_INVALID_INPUT;
		ip += csize;
		done += want;
	}
	return ip == n ? ORACLE_OK : ORACLE_INVALID_INPUT;
}
