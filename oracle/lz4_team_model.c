/*
 * TEST INFRASTRUCTURE — CPU model of b200comp's own LZ4 encoder for chunks <= 64 KiB
 * (hipcomp-core_b200/csrc/lz4_team.cuh).  It is NOT a restatement of the reference: the
 * reference's greedy one-warp encoder (src/LZ4Kernels.hiph:793-969) is restated in
 * lz4_oracle.c.  This file states, sequentially, the policy the CUDA kernel implements in
 * parallel, so that tests can demand byte-identical output from the GPU (any divergence is a
 * kernel bug) while format parity is checked separately with liblz4 / the reference decoder
 * and ratio parity against lz4_oracle_compress.
 *
 * Policy (every number below is mirrored in lz4_team.cuh):
 *   - all positions p <= n - 12 get a candidate distance d[p] (0 = none).  A position whose 4 bytes
 *     equal the 4 bytes 4 (else 8, else 1) positions earlier takes that distance (runs and repeated 1/2/4/8-byte
 *     elements never touch the table).  Every other position is hashed,
 *     h = (le32(p) * 2654435761) >> (32 - hash_log), and takes the table entry of its hash as the
 *     table stood before its aligned window of 32 positions; after the window the table holds, per
 *     hash, the highest such position of the window.
 *   - the chunk is cut into segments of 1024 positions.  Inside a segment the parse is
 *     greedy over 32-position windows that start at the cursor: the first position of the
 *     window whose candidate's 4 bytes match starts a match, extended up to
 *     E = min(n - 5, segment end); positions of the same window behind the match end are
 *     tried next, then the window moves to max(match end, window start + 32).
 *   - a match that ends exactly at a segment end stays "open": if the next segment's first
 *     sequence is a match at its first byte and the open match's offset reproduces all of it
 *     (trivially so when the offsets are equal), the two are one match with the open offset.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define SEG 1024u

static uint32_t ld32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }

static size_t put_lsic(uint8_t* out, size_t pos, uint32_t x)
{
  while (x >= 255) { out[pos++] = 0xff; x -= 255; }
  out[pos++] = (uint8_t)x;
  return pos;
}

typedef struct {
  uint8_t* out;
  size_t c;          /* bytes written */
  int open;          /* an unfinished match is pending */
  size_t tok_pos;    /* where its token goes */
  uint32_t tok_hi;   /* literal nibble of that token */
  uint32_t mlen, d;  /* its length so far and offset */
  uint32_t lit_from; /* first literal not yet written (when !open) */
} emit_t;

static void close_open(emit_t* e)
{
  const uint32_t m = e->mlen - 4;
  e->out[e->tok_pos] = (uint8_t)((e->tok_hi << 4) | (m >= 15 ? 15 : m));
  if (m >= 15) e->c = put_lsic(e->out, e->c, m - 15);
  e->open = 0;
}

size_t lz4_team_model_compress(const uint8_t* in, size_t n_, uint8_t* out, int hash_log)
{
  static uint16_t table[1 << 16];
  static uint16_t d[65536];
  const uint32_t n = (uint32_t)n_;
  if (n == 0 || n > 65536) return 0;
  memset(table, 0xff, sizeof(uint16_t) << hash_log);
  const uint32_t npos = n >= 12 ? n - 11 : 0;
  const uint32_t mlimit = n >= 5 ? n - 5 : 0;
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
        d[p] = told[l] < p ? (uint16_t)(p - told[l]) : 0;
        table[h[l]] = (uint16_t)p; /* ... and leaves its highest position per hash */
      }
    }
  }
  emit_t e;
  memset(&e, 0, sizeof e);
  e.out = out;
  for (uint32_t S0 = 0; S0 < n; S0 += SEG) {
    const uint32_t S1 = S0 + SEG < n ? S0 + SEG : n;
    const uint32_t E = mlimit < S1 ? mlimit : S1;
    const uint32_t pend = S1 < npos ? S1 : npos;
    uint32_t c = S0;
    int first = 1;
    while (c < pend) {
      uint32_t vm = 0;
      for (uint32_t l = 0; l < 32; ++l) {
        const uint32_t p = c + l;
        if (p < pend && d[p] && p + 4 <= E && ld32(in + p) == ld32(in + p - d[p])) vm |= 1u << l;
      }
      if (!vm) { c += 32; continue; }
      const uint32_t wbase = c;
      uint32_t mend = 0;
      while (vm) {
        const uint32_t ql = (uint32_t)__builtin_ctz(vm), q = wbase + ql, dq = d[q];
        uint32_t m = q + 4;
        while (m < E && in[m] == in[m - dq]) ++m;
        mend = m;
        const uint32_t mlen = mend - q;
        /* ---- emission, with the open-match rule at the segment start */
        int merged = 0;
        if (first && e.open) {
          int same = q == S0;
          if (same && dq != e.d) /* the pending match's own offset must cover this whole match */
            for (uint32_t i = 0; i < mlen && same; ++i) same = in[S0 + i] == in[S0 + i - e.d];
          if (same) {
            e.mlen += mlen;
            merged = 1;
          } else {
            close_open(&e);
            e.lit_from = S0;
          }
        }
        first = 0;
        if (!merged) {
          const uint32_t nlit = q - e.lit_from;
          e.tok_pos = e.c++;
          e.tok_hi = nlit >= 15 ? 15 : nlit;
          if (nlit >= 15) e.c = put_lsic(e.out, e.c, nlit - 15);
          memcpy(e.out + e.c, in + e.lit_from, nlit);
          e.c += nlit;
          e.out[e.c++] = (uint8_t)(dq & 0xff);
          e.out[e.c++] = (uint8_t)(dq >> 8);
          e.mlen = mlen;
          e.d = dq;
          e.open = 1;
        }
        if (mend != S1) close_open(&e); /* a match that stops at the segment end stays open */
        e.lit_from = mend;
        const uint32_t rel = mend - wbase;
        if (rel >= 32) break;
        vm &= 0xFFFFFFFFu << rel;
      }
      c = mend > wbase + 32 ? mend : wbase + 32;
    }
    if (first && e.open) { /* no sequence in this segment: the pending match ends here */
      close_open(&e);
      e.lit_from = S0;
    }
  }
  if (e.open) close_open(&e); /* cannot happen (E < n in the last segment); kept for safety */
  const uint32_t nlit = n - e.lit_from;
  e.out[e.c++] = (uint8_t)((nlit >= 15 ? 15 : nlit) << 4);
  if (nlit >= 15) e.c = put_lsic(e.out, e.c, nlit - 15);
  memcpy(e.out + e.c, in + e.lit_from, nlit);
  e.c += nlit;
  return e.c;
}

void lz4_team_model_compress_batch(const uint8_t* in, const size_t* in_off, const size_t* in_bytes, size_t batch,
                                   uint8_t* out, const size_t* out_off, size_t* out_bytes, int hash_log)
{
  for (size_t i = 0; i < batch; ++i)
    out_bytes[i] = lz4_team_model_compress(in + in_off[i], in_bytes[i], out + out_off[i], hash_log);
}
