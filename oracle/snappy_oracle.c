/*
 * TEST INFRASTRUCTURE — CPU oracle for the batched Snappy raw-block codec.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use this file, as the checker.  The product never
 * links or calls it.
 *
 * snappy_oracle_compress restates the reference encoder policy sequentially,
 * lane by lane, so its output is the byte stream the reference GPU encoder
 * produces for the same input:
 *   preamble varint        <- src/snappy/compression.hiph:316-322
 *   hash                   <- :57-60   (12 bits of v * 0x102A6B)
 *   32-position window scan, nearest-lower-lane twin, table candidate,
 *   table update           <- :190-246 (FindFourByteMatch)
 *   extension to <= 64     <- :252-269, 367-373 (Match60)
 *   literal / copy elements<- :73-151  (StoreLiterals, StoreCopy)
 *   round structure        <- :338-378
 * snappy_oracle_decompress follows the element grammar of the reference's
 * single-thread decoder (src/snappy/decompression_decode.hiph:72-150), the
 * preamble rules of src/snappy/decompression.hiph:71-104,148-158 and the
 * size query of src/lowlevel/SnappyBatchKernels.hip:84-134.
 *
 * Parity pinning (tests/test_oracle_snappy.py): the two golden streams of the
 * reference (tests/test_snappy_app.cpp:202-244, committed as
 * tests/golden/snappy_app_*.bin), the byte-exact encoder expectations and the
 * unusual-but-legal decoder streams of src/test/SnappyLargeTokens_test.cpp:381-537,
 * and cross-decoding against stock snappy (pyarrow's bundled codec).
 *
 * Differences from the reference decoder, all on malformed input only: copy
 * offsets are checked against the bytes produced *before* the element (the
 * reference compares after adding the element length, letting a copy reach up
 * to `len` bytes in front of the output buffer), and element headers/literals
 * that overrun the compressed input are errors instead of reads past the end.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define ST_SUCCESS 0
#define ST_CANNOT_DECOMPRESS 12

#define HASH_BITS 12
#define MAX_LITERAL_LENGTH 256u
#define MAX_COPY_DISTANCE 32768u

size_t snappy_oracle_max_compressed(size_t n) { return 32 + n + n / 6; }

static uint32_t snap_hash(uint32_t v) { return (v * 0x102A6Bu) >> (32 - HASH_BITS); }
static uint32_t ld32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }

static size_t put_literal(uint8_t* dst, size_t o, const uint8_t* src, uint32_t len)
{
  const uint32_t l1 = len - 1;
  if (l1 < 60) {
    dst[o++] = (uint8_t)(l1 << 2);
  } else if (l1 <= 0xff) {
    dst[o++] = 60 << 2;
    dst[o++] = (uint8_t)l1;
  } else if (l1 <= 0xffff) {
    dst[o++] = 61 << 2;
    dst[o++] = (uint8_t)l1;
    dst[o++] = (uint8_t)(l1 >> 8);
  } else if (l1 <= 0xffffff) {
    dst[o++] = 62 << 2;
    dst[o++] = (uint8_t)l1;
    dst[o++] = (uint8_t)(l1 >> 8);
    dst[o++] = (uint8_t)(l1 >> 16);
  } else {
    dst[o++] = 63 << 2;
    dst[o++] = (uint8_t)l1;
    dst[o++] = (uint8_t)(l1 >> 8);
    dst[o++] = (uint8_t)(l1 >> 16);
    dst[o++] = (uint8_t)(l1 >> 24);
  }
  memcpy(dst + o, src, len);
  return o + len;
}

static size_t put_copy(uint8_t* dst, size_t o, uint32_t len, uint32_t dist)
{
  if (len < 12 && dist < 2048) {
    dst[o++] = (uint8_t)(((dist & 0x700) >> 3) | ((len - 4) << 2) | 0x01);
    dst[o++] = (uint8_t)dist;
  } else {
    dst[o++] = (uint8_t)(((len - 1) << 2) | 0x2);
    dst[o++] = (uint8_t)dist;
    dst[o++] = (uint8_t)(dist >> 8);
  }
  return o;
}

/*
 * One search round starting at pos0: returns the literal length and sets
 * *copy_len (0 or 4) and *dist.  Mirrors the 32-lane window loop.
 */
static uint32_t find_match(const uint8_t* src, uint32_t len, uint32_t pos0, uint16_t* map, uint32_t* copy_len,
                           uint32_t* dist)
{
  uint32_t pos = pos0;
  const uint32_t maxpos = pos0 + MAX_LITERAL_LENGTH - 31;
  uint32_t literal_cnt;
  *copy_len = 0;
  do {
    uint32_t data[32], hash[32], offs[32];
    int valid[32], match[32];
    for (uint32_t t = 0; t < 32; ++t) {
      valid[t] = (pos + t + 4 <= len);
      data[t] = valid[t] ? ld32(src + pos + t) : 0;
      hash[t] = valid[t] ? snap_hash(data[t]) : 0;
    }
    for (uint32_t t = 0; t < 32; ++t) {
      match[t] = 0;
      offs[t] = pos + t;
      if (!valid[t]) continue;
      int lower = -1; /* nearest lower lane with the same hash */
      for (int u = (int)t - 1; u >= 0; --u)
        if (hash[u] == hash[t]) { lower = u; break; }
      if (lower >= 0 && data[lower] == data[t]) {
        match[t] = 1;
        offs[t] = pos + (uint32_t)lower;
      } else {
        uint32_t o = (pos & ~0xffffu) | map[hash[t]];
        if (o >= pos) o = (o >= 0x10000) ? o - 0x10000 : pos;
        offs[t] = o;
        match[t] = (o < pos && o + MAX_COPY_DISTANCE >= pos + t && ld32(src + o) == data[t]);
      }
    }
    literal_cnt = 32;
    for (uint32_t t = 0; t < 32; ++t)
      if (match[t]) { literal_cnt = t; break; }
    if (literal_cnt < 32) {
      *dist = pos + literal_cnt - offs[literal_cnt];
      *copy_len = 4;
    }
    /* table update: within lanes 0..literal_cnt, the highest lane of each
     * equal-hash group stores its position.  Lanes without 4 readable bytes
     * take part in the grouping with hash 0 but never store (:211-242). */
    const uint32_t top = literal_cnt < 32 ? literal_cnt : 31;
    for (uint32_t t = 0; t <= top; ++t) {
      if (!valid[t]) continue;
      int highest = 1;
      for (uint32_t u = t + 1; u <= top; ++u)
        if (hash[u] == hash[t]) { highest = 0; break; }
      if (highest) map[hash[t]] = (uint16_t)(pos + t);
    }
    pos += literal_cnt;
  } while (literal_cnt == 32 && pos < maxpos);
  return (pos < len ? pos : len) - pos0;
}

size_t snappy_oracle_compress(const uint8_t* src, size_t src_len, uint8_t* dst)
{
  uint16_t map[1 << HASH_BITS];
  memset(map, 0, sizeof(map));
  const uint32_t len = (uint32_t)src_len;
  size_t o = 0;
  uint32_t v = len;
  while (v > 0x7f) { dst[o++] = (uint8_t)(v | 0x80); v >>= 7; }
  dst[o++] = (uint8_t)v;
  uint32_t pos = 0;
  while (pos < len) {
    uint32_t copy_len, dist = 0;
    const uint32_t lit = find_match(src, len, pos, map, &copy_len, &dist);
    if (copy_len) {
      const uint32_t mpos = pos + lit + 4;
      uint32_t lim = len - mpos;
      if (lim > 60) lim = 60;
      uint32_t k = 0;
      while (k < lim && src[mpos + k] == src[mpos - dist + k]) ++k;
      copy_len += k;
    }
    if (lit) o = put_literal(dst, o, src + pos, lit);
    pos += lit;
    if (copy_len) o = put_copy(dst, o, copy_len, dist);
    pos += copy_len;
  }
  return o;
}

/* varint preamble as read by the size query (value 0 on >= 2^31) */
size_t snappy_oracle_get_size(const uint8_t* comp, size_t comp_len)
{
  if (comp_len == 0) return 0;
  uint32_t size = 0;
  size_t cur = 0;
  for (int shift = 0; shift <= 28; shift += 7) {
    const uint32_t c = cur < comp_len ? comp[cur++] : 0;
    if (shift == 28) {
      if (c >= 8) return 0;
      size |= c << 28;
      break;
    }
    size |= (c & 0x7f) << shift;
    if (!(c & 0x80)) break;
  }
  return size;
}

int snappy_oracle_decompress(const uint8_t* comp, size_t comp_len, uint8_t* out, size_t cap, size_t* actual)
{
  *actual = 0;
  if (comp_len == 0) return ST_CANNOT_DECOMPRESS;
  uint32_t size = 0;
  size_t cur = 0;
  for (int shift = 0; shift <= 28; shift += 7) {
    const uint32_t c = cur < comp_len ? comp[cur++] : 0;
    if (shift == 28) {
      if (c >= 8) return ST_CANNOT_DECOMPRESS;
      size |= c << 28;
      break;
    }
    size |= (c & 0x7f) << shift;
    if (!(c & 0x80)) break;
  }
  if (cap == 0) cap = size; /* reference: capacity 0 means "trust the preamble" */
  if ((cur >= comp_len && size != 0) || size > cap) return ST_CANNOT_DECOMPRESS;
  size_t left = size, dst = 0;
  int ok = 1;
  while (left > 0) {
    if (cur >= comp_len) { ok = 0; break; }
    const uint8_t b0 = comp[cur];
    uint32_t blen, offset = 0;
    if (b0 & 3) {
      if ((b0 & 3) == 1) {
        if (cur + 2 > comp_len) { ok = 0; break; }
        offset = ((uint32_t)(b0 & 0xe0) << 3) | comp[cur + 1];
        blen = ((b0 >> 2) & 7) + 4;
        cur += 2;
      } else if ((b0 & 3) == 2) {
        if (cur + 3 > comp_len) { ok = 0; break; }
        offset = comp[cur + 1] | ((uint32_t)comp[cur + 2] << 8);
        blen = (b0 >> 2) + 1;
        cur += 3;
      } else {
        if (cur + 5 > comp_len) { ok = 0; break; }
        offset = ld32(comp + cur + 1);
        blen = (b0 >> 2) + 1;
        cur += 5;
      }
      if (offset == 0 || offset > dst || left < blen) { ok = 0; break; }
      for (uint32_t i = 0; i < blen; ++i) out[dst + i] = out[dst + i - offset];
    } else {
      blen = b0 >> 2;
      cur += 1;
      if (blen >= 60) {
        const uint32_t nb = blen - 59;
        if (cur + nb > comp_len) { ok = 0; break; }
        blen = 0;
        for (uint32_t i = 0; i < nb; ++i) blen |= (uint32_t)comp[cur + i] << (8 * i);
        cur += nb;
      }
      if (blen == 0xffffffffu) { ok = 0; break; }
      blen += 1;
      if (left < blen || cur + blen > comp_len) { ok = 0; break; }
      memcpy(out + dst, comp + cur, blen);
      cur += blen;
    }
    dst += blen;
    left -= blen;
  }
  *actual = size - left;
  return ok ? ST_SUCCESS : ST_CANNOT_DECOMPRESS;
}

void snappy_oracle_compress_batch(const uint8_t* in, const size_t* in_off, const size_t* in_bytes, size_t batch,
                                  uint8_t* out, const size_t* out_off, size_t* out_bytes)
{
  for (size_t i = 0; i < batch; ++i)
    out_bytes[i] = snappy_oracle_compress(in + in_off[i], in_bytes[i], out + out_off[i]);
}

void snappy_oracle_decompress_batch(const uint8_t* comp, const size_t* comp_off, const size_t* comp_bytes,
                                    size_t batch, uint8_t* out, const size_t* out_off, const size_t* out_cap,
                                    size_t* actual, int* status)
{
  for (size_t i = 0; i < batch; ++i)
    status[i] = snappy_oracle_decompress(comp + comp_off[i], comp_bytes[i], out + out_off[i], out_cap[i], &actual[i]);
}
