/*
 * TEST INFRASTRUCTURE — CPU oracle for the batched LZ4 block codec.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use this file, as the checker.  The product never
 * links or calls it.
 *
 * lz4_oracle_compress is a sequential restatement of the reference's
 * one-warp-per-chunk greedy encoder, lane by lane, so that its output is the
 * byte stream the reference encoder produces for the same input, data type
 * and hash-table size:
 *   hash              <- src/LZ4Kernels.hiph:557-561
 *   table size        <- src/lowlevel/LZ4CompressionKernels.hip:142-156
 *   candidate search  <- src/LZ4Kernels.hiph:829-923 (window keys :743-791,
 *                        intra-window twins :867-894, table probe :634-663,
 *                        convertIdx :619-632)
 *   table insertion   <- :722-741 (highest lane wins a slot)
 *   match extension   <- :592-617
 *   sequence emission <- :665-715, token nibbles :280-351, LSIC :267-278
 * lz4_oracle_decompress follows decompressStream (:971-1097) including its
 * three out-of-bounds conditions, and additionally bounds every compressed
 * read (the reference can read past a truncated chunk).
 *
 * Parity pinning: the reference holds no byte-level LZ4 vectors,
 * so the format is pinned at the third-party boundary it names — the LZ4 block
 * format as implemented by liblz4 1.9.4 (/lib/x86_64-linux-gnu/liblz4.so.1):
 * tests/test_oracle_lz4.py decodes this encoder's output with
 * LZ4_decompress_safe and feeds LZ4_compress_default output to this decoder.
 * Encoder *policy* parity with the reference GPU encoder is "unpinned" until
 * the reference build (oracle/_ref) has run on a GPU box; see DESIGN.md.
 *
 * One deliberate difference: for 2- and 4-byte data types the reference
 * stores (distance_in_elements * sizeof(T)) in a 16-bit field (:935,:954), which
 * silently truncates distances >= 65536 bytes in chunks larger than 64 KiB.
 * Here such candidates are rejected.  For chunks <= 64 KiB the two agree.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define ST_SUCCESS 0
#define ST_CANNOT_DECOMPRESS 12
#define NULL_OFFSET 0xFFFFu
#define MAX_OFFSET 65535u

static uint32_t brev32(uint32_t v)
{
  v = ((v >> 1) & 0x55555555u) | ((v & 0x55555555u) << 1);
  v = ((v >> 2) & 0x33333333u) | ((v & 0x33333333u) << 2);
  v = ((v >> 4) & 0x0F0F0F0Fu) | ((v & 0x0F0F0F0Fu) << 4);
  v = ((v >> 8) & 0x00FF00FFu) | ((v & 0x00FF00FFu) << 8);
  return (v >> 16) | (v << 16);
}
static uint32_t lz4_hash(uint32_t key, uint32_t ht) { return (brev32(key) + (key ^ 0xc375u)) & (ht - 1); }
static uint32_t ld32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }

uint32_t lz4_oracle_hash_table_size(size_t max_chunk)
{
  size_t p = 1;
  while (p < max_chunk) p *= 2;
  return (uint32_t)(p < 16384 ? p : 16384);
}

size_t lz4_oracle_max_compressed(size_t n) { return (n + 1 + (n + 254) / 255 + 7) / 8 * 8; }

static size_t put_lsic(uint8_t* out, size_t pos, uint32_t x)
{
  const uint32_t num = x / 255 + 1;
  for (uint32_t i = 0; i + 1 < num; ++i) out[pos++] = 0xff;
  out[pos++] = (uint8_t)(x % 255);
  return pos;
}

static size_t put_sequence(uint8_t* out, size_t cpos, const uint8_t* lit, uint32_t nlit, uint32_t nmatch,
                           uint32_t offset)
{
  const uint8_t hi = (uint8_t)(nlit >= 15 ? 15 : nlit);
  /* the reference computes (num_matches - 4) in uint8 for the literal-only
   * last token too, leaving 0xC in the low nibble (:324-331) */
  const uint8_t lo = (uint8_t)(nmatch >= 19 ? 15 : ((nmatch - 4) & 0x0f));
  out[cpos++] = (uint8_t)((hi << 4) | lo);
  if (nlit >= 15) cpos = put_lsic(out, cpos, nlit - 15);
  memcpy(out + cpos, lit, nlit);
  cpos += nlit;
  if (nmatch > 0) {
    out[cpos++] = (uint8_t)(offset & 0xff);
    out[cpos++] = (uint8_t)(offset >> 8);
    if (nmatch >= 19) cpos = put_lsic(out, cpos, nmatch - 19);
  }
  return cpos;
}

/*
 * tsize: 1, 2 or 4 (match granularity = data_type width).
 * ht_size: entries in the hash table (power of two, <= 16384).
 * Returns the compressed size.
 */
size_t lz4_oracle_compress(const uint8_t* in, size_t length, uint8_t* out, int tsize, uint32_t ht_size)
{
  static const uint32_t W = 32;
  uint16_t table[16384];
  for (uint32_t i = 0; i < ht_size; ++i) table[i] = NULL_OFFSET;
  const uint32_t sT = (uint32_t)tsize;
  const uint32_t typed_length = (uint32_t)((length + sT - 1) / sT);
  const uint32_t last_valid_match = (12 + sT - 1) / sT;
  const uint32_t min_ending = (5 + sT - 1) / sT;
  const uint32_t invalid_threads = 3 / sT;
  uint32_t d = 0;
  size_t c = 0;
  while (d < typed_length) {
    const uint32_t token_start = d;
    for (;;) {
      if (d + last_valid_match >= typed_length) {
        d = typed_length;
        c = put_sequence(out, c, in + (size_t)token_start * sT, (uint32_t)(length - (size_t)token_start * sT), 0, 0);
        break;
      }
      uint32_t nvalid = W - invalid_threads;
      if (typed_length - d - last_valid_match < nvalid) nvalid = typed_length - d - last_valid_match;
      uint32_t key[32];
      for (uint32_t t = 0; t < nvalid; ++t) key[t] = ld32(in + (size_t)(d + t) * sT);
      /* intra-window twins: first lane that has an earlier lane with its key */
      uint32_t first = nvalid, match_loc = typed_length;
      for (uint32_t t = 1; t < nvalid && first == nvalid; ++t)
        for (uint32_t u = 0; u < t; ++u)
          if (key[u] == key[t]) { first = t; match_loc = d + u; break; }
      /* table probe by the lanes in front of it */
      for (uint32_t t = 0; t < first; ++t) {
        const uint16_t h = table[lz4_hash(key[t], ht_size)];
        if (h == NULL_OFFSET) continue;
        const uint32_t pos = d + t;
        uint32_t cand = (pos / 65536u) * 65536u + h;
        if (cand >= pos) cand -= 65536u;
        if (pos - cand > MAX_OFFSET) continue;
        if ((uint64_t)(pos - cand) * sT > MAX_OFFSET) continue; /* see header note */
        if (ld32(in + (size_t)cand * sT) != key[t]) continue;
        first = t;
        match_loc = cand;
        break;
      }
      if (match_loc != typed_length) {
        for (uint32_t t = 0; t < first; ++t) table[lz4_hash(key[t], ht_size)] = (uint16_t)((d + t) & MAX_OFFSET);
        const uint32_t pos = d + first;
        const uint32_t limit = typed_length - pos - min_ending;
        uint32_t mlen = 0;
        while (mlen < limit
               && memcmp(in + (size_t)(match_loc + mlen) * sT, in + (size_t)(pos + mlen) * sT, sT) == 0)
          ++mlen;
        const uint32_t nlit = pos - token_start;
        d = token_start + nlit + mlen;
        c = put_sequence(out, c, in + (size_t)token_start * sT, nlit * sT, mlen * sT, (pos - match_loc) * sT);
        break;
      }
      for (uint32_t t = 0; t < nvalid; ++t) table[lz4_hash(key[t], ht_size)] = (uint16_t)((d + t) & MAX_OFFSET);
      d += nvalid;
    }
  }
  return c;
}

/*
 * Decode one LZ4 block.  write_output == 0 is the size-only walk
 * (capacity unbounded).  Returns status; *actual = decoded size or 0.
 */
int lz4_oracle_decompress(const uint8_t* comp, size_t comp_len, uint8_t* out, size_t cap, int write_output,
                          size_t* actual)
{
  size_t ci = 0, di = 0;
  *actual = 0;
  if (!write_output) cap = (size_t)0xFFFFFFFFu;
  while (ci < comp_len) {
    const uint8_t tok = comp[ci++];
    size_t nlit = tok >> 4;
    if (nlit == 15) {
      uint8_t b = 0xff;
      while (b == 0xff) {
        if (ci >= comp_len) return ST_CANNOT_DECOMPRESS;
        b = comp[ci++];
        nlit += b;
      }
    }
    if (di + nlit > cap) return ST_CANNOT_DECOMPRESS;
    if (ci + nlit > comp_len) return ST_CANNOT_DECOMPRESS;
    if (write_output) memcpy(out + di, comp + ci, nlit);
    ci += nlit;
    di += nlit;
    if (ci < comp_len) {
      if (ci + 2 > comp_len) return ST_CANNOT_DECOMPRESS;
      const size_t offset = (size_t)comp[ci] | ((size_t)comp[ci + 1] << 8);
      ci += 2;
      size_t mlen = 4 + (tok & 15);
      if ((tok & 15) == 15) {
        uint8_t b = 0xff;
        while (b == 0xff) {
          if (ci >= comp_len) return ST_CANNOT_DECOMPRESS;
          b = comp[ci++];
          mlen += b;
        }
      }
      if (offset == 0 || di < offset || di + mlen > cap) return ST_CANNOT_DECOMPRESS;
      if (write_output)
        for (size_t i = 0; i < mlen; ++i) out[di + i] = out[di + i - offset];
      di += mlen;
    }
  }
  *actual = di;
  return ST_SUCCESS;
}

void lz4_oracle_compress_batch(const uint8_t* in, const size_t* in_off, const size_t* in_bytes, size_t batch,
                               uint8_t* out, const size_t* out_off, size_t* out_bytes, int tsize, uint32_t ht_size)
{
  for (size_t i = 0; i < batch; ++i)
    out_bytes[i] = lz4_oracle_compress(in + in_off[i], in_bytes[i], out + out_off[i], tsize, ht_size);
}

void lz4_oracle_decompress_batch(const uint8_t* comp, const size_t* comp_off, const size_t* comp_bytes, size_t batch,
                                 uint8_t* out, const size_t* out_off, const size_t* out_cap, size_t* actual,
                                 int* status)
{
  for (size_t i = 0; i < batch; ++i)
    status[i] = lz4_oracle_decompress(comp + comp_off[i], comp_bytes[i], out + out_off[i], out_cap[i], 1, &actual[i]);
}
