/*
 * TEST INFRASTRUCTURE — CPU model of b200comp's own LZ4 and Snappy encoders for chunks <= 64 KiB
 * (hipcomp-core_b200/csrc/lz_team_encode.cuh; the two codecs share the match finder and differ in how a
 * sequence is written).  It is NOT a restatement of the reference: the
 * reference's greedy encoders (src/LZ4Kernels.hiph:793-969, src/snappy/compression.hiph:190-385) are
 * restated in lz4_oracle.c / snappy_oracle.c.  This file states, sequentially, the policy the CUDA kernel implements in
 * parallel, so that tests can demand byte-identical output from the GPU (any divergence is a
 * kernel bug) while format parity is checked separately with liblz4 / the reference decoder
 * and ratio parity against lz4_oracle_compress.
 *
 * Policy (every number below is mirrored in lz_team_encode.cuh):
 *   1. candidates.  Every position p <= n - 12 gets a candidate distance d[p] (0 = none).  A position whose
 *      4 bytes equal the 4 bytes 4 (else 8, else 1) positions earlier takes that distance (runs and repeated
 *      1/2/4/8-byte elements never touch the table).  Every other position is hashed,
 *      h = (le32(p) * 2654435761) >> (32 - hash_log), and takes the table entry of its hash as the table
 *      stood before its aligned window of 32 positions; after the window the table holds, per hash, the
 *      highest position inserted so far.
 *   2. block parse.  Every aligned block of 32 positions is parsed greedily on its own, starting at its first
 *      position: a position whose candidate's 4 bytes match starts a match, which is extended (beyond the
 *      block) up to n - 5; the parse continues behind it while it is inside the block.  A match still equal
 *      after 4 + 4*LANE_STEPS bytes is "long": its end is settled in step 3.
 *   3. selection, in position order, with X = the largest match end seen so far (0 at the start):
 *      a long match whose verified part ends at or before X is dropped; any other long match is extended
 *      to its true end.  A match [q, e) is cut to [max(q, X), e) and emitted when at least 4 bytes remain
 *      (otherwise dropped); either way X = max(X, e).  Literals run from the end of the last emitted match.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define BLK 32u
#define LANE_STEPS 15u

static uint32_t ld32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }

static size_t put_lsic(uint8_t* out, size_t pos, uint32_t x)
{
  while (x >= 255) { out[pos++] = 0xff; x -= 255; }
  out[pos++] = (uint8_t)x;
  return pos;
}

/* a sequence: literals [L, s) followed by a match of length m - s at distance dq (m == s: the final literals) */
typedef size_t (*emit_fn)(uint8_t* out, size_t c, const uint8_t* in, uint32_t L, uint32_t s, uint32_t m, uint32_t dq);

static size_t lz4_emit(uint8_t* out, size_t c, const uint8_t* in, uint32_t L, uint32_t s, uint32_t m, uint32_t dq)
{
  const uint32_t nlit = s - L;
  if (m == s) { /* last sequence: literals only */
    out[c++] = (uint8_t)((nlit >= 15 ? 15 : nlit) << 4);
    if (nlit >= 15) c = put_lsic(out, c, nlit - 15);
    memcpy(out + c, in + L, nlit);
    return c + nlit;
  }
  const uint32_t ml = m - s - 4;
  out[c++] = (uint8_t)(((nlit >= 15 ? 15 : nlit) << 4) | (ml >= 15 ? 15 : ml));
  if (nlit >= 15) c = put_lsic(out, c, nlit - 15);
  memcpy(out + c, in + L, nlit);
  c += nlit;
  out[c++] = (uint8_t)(dq & 0xff);
  out[c++] = (uint8_t)(dq >> 8);
  if (ml >= 15) c = put_lsic(out, c, ml - 15);
  return c;
}

/* Snappy elements (format: share/snappy/hipcomp_snappy.py:113-181, src/snappy/symbol.hiph:43-70): one literal
 * element when there are literals, then copies of at most 64 bytes; the 2-byte copy form when it fits. */
static size_t snappy_emit(uint8_t* out, size_t c, const uint8_t* in, uint32_t L, uint32_t s, uint32_t m, uint32_t dq)
{
  const uint32_t nlit = s - L;
  if (nlit) {
    const uint32_t l1 = nlit - 1;
    if (l1 < 60) {
      out[c++] = (uint8_t)(l1 << 2);
    } else if (l1 <= 0xff) {
      out[c++] = 60 << 2;
      out[c++] = (uint8_t)l1;
    } else {
      out[c++] = 61 << 2;
      out[c++] = (uint8_t)l1;
      out[c++] = (uint8_t)(l1 >> 8);
    }
    memcpy(out + c, in + L, nlit);
    c += nlit;
  }
  uint32_t len = m - s;
  while (len) {
    uint32_t piece = len;
    if (len >= 68) piece = 64;
    else if (len > 64) piece = 60; /* leaves at least 4 for the last copy */
    if (piece < 12 && dq < 2048) {
      out[c++] = (uint8_t)(((dq & 0x700) >> 3) | ((piece - 4) << 2) | 0x01);
      out[c++] = (uint8_t)dq;
    } else {
      out[c++] = (uint8_t)(((piece - 1) << 2) | 0x2);
      out[c++] = (uint8_t)dq;
      out[c++] = (uint8_t)(dq >> 8);
    }
    len -= piece;
  }
  return c;
}

static size_t team_model(const uint8_t* in, uint32_t n, uint8_t* out, size_t c, int hash_log, emit_fn emit)
{
  static uint32_t table[1 << 16]; /* position + 1; 0 = empty */
  static uint16_t d[65536];
  memset(table, 0, sizeof(uint32_t) << hash_log);
  const uint32_t npos = n >= 12 ? n - 11 : 0;
  const uint32_t E = n >= 5 ? n - 5 : 0; /* matches end at or before n - 5 */
  for (uint32_t w0 = 0; w0 < npos; w0 += 32) {
    uint32_t h[32], told[32];
    const uint32_t cnt = npos - w0 < 32 ? npos - w0 : 32;
    for (uint32_t l = 0; l < cnt; ++l) {
      const uint32_t p = w0 + l, key = ld32(in + p);
      if (p >= 4 && key == ld32(in + p - 4)) h[l] = 0x8004u;      /* settled without the table */
      else if (p >= 8 && key == ld32(in + p - 8)) h[l] = 0x8008u;
      else if (p >= 1 && key == ld32(in + p - 1)) h[l] = 0x8001u;
      else h[l] = (key * 2654435761u) >> (32 - hash_log);
    }
    /* the window reads the table as the previous windows left it ... */
    for (uint32_t l = 0; l < cnt; ++l) told[l] = (h[l] & 0x8000u) ? 0 : table[h[l]];
    for (uint32_t l = 0; l < cnt; ++l) {
      const uint32_t p = w0 + l;
      if (h[l] & 0x8000u) {
        d[p] = (uint16_t)(h[l] & 0x7fffu);
      } else {
        d[p] = told[l] ? (uint16_t)(p + 1 - told[l]) : 0;
        table[h[l]] = p + 1; /* ... and leaves its highest position per hash */
      }
    }
  }
  uint32_t X = 0, L = 0; /* largest match end seen; first literal not yet written */
  for (uint32_t b0 = 0; b0 < npos; b0 += BLK) {
    const uint32_t bend = b0 + BLK < npos ? b0 + BLK : npos;
    uint32_t p = b0;
    while (p < bend) {
      const uint32_t dq = d[p];
      if (!dq || ld32(in + p) != ld32(in + p - dq)) { ++p; continue; }
      const uint32_t q = p;
      /* the owner of the block verifies up to LANE_STEPS further words */
      uint32_t m = q + 4, steps = 0;
      int is_long = 0;
      for (;;) {
        if (m >= E) { m = E; break; }
        if (steps == LANE_STEPS) { is_long = 1; break; }
        const uint32_t x = ld32(in + m) ^ ld32(in + m - dq);
        if (x) { m += (uint32_t)__builtin_ctz(x) >> 3; if (m > E) m = E; break; }
        m += 4; ++steps;
      }
      if (is_long && m > X) { /* the team settles the true end */
        while (m < E && in[m] == in[m - dq]) ++m;
      }
      const uint32_t s = q > X ? q : X;
      if (m >= s + 4) {
        c = emit(out, c, in, L, s, m, dq);
        L = m;
      }
      if (m > X) X = m;
      p = m; /* m >= q + 4 */
    }
  }
  return emit(out, c, in, L, n, n, 0);
}

size_t lz4_team_model_compress(const uint8_t* in, size_t n, uint8_t* out, int hash_log)
{
  if (n == 0 || n > 65536) return 0;
  return team_model(in, (uint32_t)n, out, 0, hash_log, lz4_emit);
}

size_t snappy_team_model_compress(const uint8_t* in, size_t n, uint8_t* out, int hash_log)
{
  if (n > 65536) return 0;
  size_t c = 0;
  uint32_t v = (uint32_t)n; /* preamble: uncompressed length as a varint */
  while (v >= 0x80) { out[c++] = (uint8_t)(v | 0x80); v >>= 7; }
  out[c++] = (uint8_t)v;
  if (n == 0) return c;
  return team_model(in, (uint32_t)n, out, c, hash_log, snappy_emit);
}

void lz4_team_model_compress_batch(const uint8_t* in, const size_t* in_off, const size_t* in_bytes, size_t batch,
                                   uint8_t* out, const size_t* out_off, size_t* out_bytes, int hash_log)
{
  for (size_t i = 0; i < batch; ++i)
    out_bytes[i] = lz4_team_model_compress(in + in_off[i], in_bytes[i], out + out_off[i], hash_log);
}
