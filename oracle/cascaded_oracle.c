/*
 * TEST INFRASTRUCTURE — CPU oracle for the batched Cascaded codec.
 *
 * Plain-C, single-threaded restatement of the reference algorithm.  It is only
 * ever used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs as the checker; the product (libhipcomp.so) never
 * links, loads or calls it.
 *
 * Parity pinning: checked against the reference's own known-answer test
 * (tests/test_cascaded_batch.cpp:212-378: header word, chunk metadata, run
 * arrays, final array, delta seed for all 8 element types), its fallback test
 * (:491-615) and its out-of-bounds matrix (:794-814) in
 * tests/test_oracle_cascaded.py.
 *
 * Follows, function by function:
 *   casc_chunk_meta_size   <- src/CascadedKernels.hiph:101-106
 *   rle_layer              <- :124-241  (block_rle_compress)
 *   delta_layer            <- :317-328  (block_delta_compress)
 *   for_bitwidth           <- :394-471  (get_for_bitwidth)
 *   write_array            <- :482-553, 646-680 (block_bitpack, block_write)
 *   casc_oracle_compress   <- :761-1058 (do_cascaded_compression_kernel)
 *   read_array             <- :563-618, 702-737 (block_bitunpack, block_read)
 *   casc_oracle_decompress <- :1106-1435 (cascaded_decompression_fcn)
 *   casc_oracle_get_size   <- src/lowlevel/CascadedBatch.hip:262-281
 *
 * Deliberate, documented differences from the reference:
 *   - alignment padding bytes, which the reference leaves uninitialised, are
 *     written as zero;
 *   - an empty bit-packed array stores frame-of-reference 0 and bit width 0
 *     (the reference stores uninitialised values there; count==0 makes them
 *     don't-care, SURVEY A.3 caveat 1);
 *   - a delta layer that meets an empty array (undefined behaviour in the
 *     reference, caveat 2) makes the partition fall back to a raw copy;
 *   - the decoder bounds every array against the 4096-byte chunk capacity
 *     instead of overrunning its scratch on corrupt input.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#define CASC_CHUNK_BYTES 4096
#define ST_SUCCESS 0
#define ST_CANNOT_DECOMPRESS 12

static size_t round_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static size_t max_sz(size_t a, size_t b) { return a > b ? a : b; }

static int type_width(int type)
{
  switch (type) {
  case 0: case 1: return 1;
  case 2: case 3: return 2;
  case 4: case 5: return 4;
  case 6: case 7: return 8;
  default: return 0;
  }
}

/* elements are kept as zero-extended uint64_t of the given width */
static uint64_t ld_elem(const uint8_t* p, int w)
{
  uint64_t v = 0;
  memcpy(&v, p, (size_t)w); /* little-endian host */
  return v;
}
static void st_elem(uint8_t* p, int w, uint64_t v) { memcpy(p, &v, (size_t)w); }
static uint64_t wmask(int w) { return w == 8 ? ~0ull : ((1ull << (8 * w)) - 1); }
static int64_t sext(uint64_t v, int w)
{
  const int sh = 64 - 8 * w;
  return (int64_t)(v << sh) >> sh;
}

size_t casc_chunk_meta_size(int w, int num_RLEs, int num_deltas)
{
  return round_up(4 + 4 * (size_t)(num_RLEs + 1), (size_t)w)
         + round_up((size_t)w * (size_t)num_deltas, 4);
}

/* run-length encode vals[0..n) in place; counts[] receives run lengths */
static size_t rle_layer(uint64_t* vals, size_t n, uint16_t* counts)
{
  size_t m = 0, start = 0;
  for (size_t i = 0; i < n; ++i) {
    if (i + 1 == n || vals[i + 1] != vals[i]) {
      counts[m] = (uint16_t)(i + 1 - start);
      vals[m] = vals[i];
      ++m;
      start = i + 1;
    }
  }
  return m;
}

/* adjacent differences in place; returns the seed (first element) */
static uint64_t delta_layer(uint64_t* vals, size_t n, int w)
{
  const uint64_t seed = vals[0];
  for (size_t i = 0; i + 1 < n; ++i)
    vals[i] = (vals[i + 1] - vals[i]) & wmask(w);
  return seed;
}

/* signed minimum (frame of reference) and bit width of max-min */
static void for_bitwidth(const uint64_t* v, size_t n, int w, uint64_t* fr, uint32_t* bw)
{
  if (n == 0) { *fr = 0; *bw = 0; return; }
  int64_t mn = sext(v[0], w), mx = mn;
  for (size_t i = 1; i < n; ++i) {
    const int64_t s = sext(v[i], w);
    if (s < mn) mn = s;
    if (s > mx) mx = s;
  }
  *fr = (uint64_t)mn & wmask(w);
  if (w > 4) {
    const uint64_t range = (uint64_t)mx - (uint64_t)mn;
    *bw = range ? (uint32_t)(64 - __builtin_clzll(range)) : 0;
  } else {
    const uint32_t range = (uint32_t)mx - (uint32_t)mn;
    *bw = range ? (uint32_t)(32 - __builtin_clz(range)) : 0;
  }
}

static size_t packed_header_bytes(int w) { return round_up((size_t)w + 4, max_sz(4, (size_t)w)); }

/*
 * Serialise v[0..n) (element width w) at out+pos, raw or bit-packed.
 * Returns 0 and sets *out_bytes, or -1 when the (4-byte padded) array would
 * pass `limit` (offsets are relative to the partition start).
 */
static int write_array(uint8_t* out, size_t pos, size_t limit, const uint64_t* v, size_t n,
                       int w, int use_bp, size_t* out_bytes)
{
  if (!use_bp) {
    const size_t bytes = n * (size_t)w;
    const size_t padded = round_up(bytes, 4);
    if (pos + padded > limit) return -1;
    for (size_t i = 0; i < n; ++i) st_elem(out + pos + i * (size_t)w, w, v[i]);
    memset(out + pos + bytes, 0, padded - bytes);
    *out_bytes = bytes;
    return 0;
  }
  uint64_t fr;
  uint32_t bw;
  for_bitwidth(v, n, w, &fr, &bw);
  const size_t words = (n * bw + 31) / 32;
  const size_t hdr = packed_header_bytes(w);
  const size_t bytes = hdr + 4 * words;
  if (pos + bytes > limit) return -1;
  uint8_t* p = out + pos;
  memset(p, 0, bytes);
  st_elem(p, w, fr);
  const uint32_t meta = (bw << 16) | (uint32_t)n;
  memcpy(p + round_up((size_t)w, 4), &meta, 4);
  uint8_t* data = p + hdr;
  for (size_t i = 0; i < n; ++i) {
    const uint64_t d = (v[i] - fr) & wmask(w);
    size_t bit = i * bw;
    for (uint32_t b = 0; b < bw; ++b, ++bit)
      if ((d >> b) & 1) data[bit >> 3] |= (uint8_t)(1u << (bit & 7));
  }
  *out_bytes = bytes;
  return 0;
}

/*
 * Compress one partition.  `out` must hold round_up(in_bytes,4)+8 bytes.
 * Returns the compressed size in bytes (0 for an empty partition).
 */
size_t casc_oracle_compress(const void* in_v, size_t in_bytes, void* out_v, int type,
                            int num_RLEs, int num_deltas, int use_bp)
{
  const uint8_t* in = (const uint8_t*)in_v;
  uint8_t* out = (uint8_t*)out_v;
  const int w = type_width(type);
  if (!in || in_bytes == 0 || w == 0) return 0;
  const size_t n_total = in_bytes / (size_t)w;
  const size_t limit = 8 + round_up(in_bytes, 4);
  const size_t chunk_elems = CASC_CHUNK_BYTES / (size_t)w;
  const size_t hdr = round_up(8, (size_t)w);
  int use_compression = !(num_RLEs == 0 && num_deltas == 0 && use_bp == 0);

  uint64_t vals[CASC_CHUNK_BYTES];
  uint64_t cnt64[CASC_CHUNK_BYTES];
  uint16_t counts[CASC_CHUNK_BYTES];
  size_t cur = hdr;
  const size_t meta_size = casc_chunk_meta_size(w, num_RLEs, num_deltas);
  if (num_RLEs < 0 || num_RLEs > 7 || num_deltas < 0 || meta_size > 64)
    use_compression = 0; /* reference asserts (:826,:1209); the library rejects these options */
  const size_t n_chunks = (n_total + chunk_elems - 1) / chunk_elems;

  for (size_t c = 0; c < n_chunks && use_compression; ++c) {
    const size_t chunk_start = cur;
    uint32_t meta[16];
    uint8_t meta_bytes[64];
    memset(meta, 0, sizeof(meta));
    memset(meta_bytes, 0, sizeof(meta_bytes));
    cur += meta_size;
    size_t n = n_total - c * chunk_elems;
    if (n > chunk_elems) n = chunk_elems;
    for (size_t i = 0; i < n; ++i) vals[i] = ld_elem(in + (c * chunk_elems + i) * (size_t)w, w);

    int rle_left = num_RLEs, delta_left = num_deltas;
    const int layers = num_RLEs > num_deltas ? num_RLEs : num_deltas;
    const size_t delta_off = round_up(4 + 4 * (size_t)(num_RLEs + 1), (size_t)w);
    for (int l = 0; l < layers && use_compression; ++l) {
      if (rle_left > 0) {
        n = rle_layer(vals, n, counts);
        for (size_t i = 0; i < n; ++i) cnt64[i] = counts[i];
        size_t ob;
        if (write_array(out, cur, limit, cnt64, n, 2, use_bp, &ob) != 0) { use_compression = 0; break; }
        cur += round_up(ob, 4);
        meta[num_RLEs - rle_left + 1] = (uint32_t)ob;
        --rle_left;
      }
      if (delta_left > 0) {
        if (n == 0) { use_compression = 0; break; } /* reference: undefined behaviour */
        const uint64_t seed = delta_layer(vals, n, w);
        st_elem(meta_bytes + delta_off + (size_t)(num_deltas - delta_left) * (size_t)w, w, seed);
        n -= 1;
        --delta_left;
      }
    }
    if (!use_compression) break;
    const size_t final_pos = round_up(cur, (size_t)w);
    size_t ob;
    if (write_array(out, final_pos, limit, vals, n, w, use_bp, &ob) != 0) { use_compression = 0; break; }
    memset(out + cur, 0, final_pos - cur);
    cur = final_pos + round_up(ob, 4);
    const size_t next = round_up(cur, (size_t)w);
    memset(out + cur, 0, next - cur);
    cur = next;
    meta[0] = (uint32_t)(cur - chunk_start);
    meta[num_RLEs + 1] = (uint32_t)ob;
    memcpy(meta_bytes, meta, 4 * (size_t)(num_RLEs + 2));
    memcpy(out + chunk_start, meta_bytes, meta_size);
  }

  size_t comp_bytes;
  memset(out, 0, hdr);
  out[3] = (uint8_t)type;
  if (use_compression) {
    out[0] = (uint8_t)num_RLEs;
    out[1] = (uint8_t)num_deltas;
    out[2] = (uint8_t)use_bp;
    comp_bytes = cur;
  } else {
    const size_t raw = n_total * (size_t)w;
    memcpy(out + hdr, in, raw);
    memset(out + hdr + raw, 0, round_up(raw, 4) - raw);
    comp_bytes = hdr + round_up(raw, 4);
  }
  const uint32_t ub = (uint32_t)(n_total * (size_t)w);
  memcpy(out + 4, &ub, 4);
  return comp_bytes;
}

/* read an array of element width w at comp+pos of `in_bytes` bytes */
static int read_array(const uint8_t* comp, size_t pos, size_t in_bytes, size_t end, int w, int use_bp,
                      uint64_t* v, size_t cap, size_t* n_out)
{
  if (pos + round_up(in_bytes, 4) > end) return -1;
  if (!use_bp) {
    const size_t n = in_bytes / (size_t)w;
    if (n > cap) return -1;
    for (size_t i = 0; i < n; ++i) v[i] = ld_elem(comp + pos + i * (size_t)w, w);
    *n_out = n;
    return 0;
  }
  const size_t hdr = packed_header_bytes(w);
  if (in_bytes < hdr) return -1;
  const uint64_t fr = ld_elem(comp + pos, w);
  uint32_t meta;
  memcpy(&meta, comp + pos + round_up((size_t)w, 4), 4);
  const uint32_t bw = meta >> 16;
  const size_t n = meta & 0xffff;
  if (n > cap || bw > (uint32_t)(8 * w)) return -1;
  if (hdr + 4 * ((n * bw + 31) / 32) > round_up(in_bytes, 4)) return -1;
  const uint8_t* data = comp + pos + hdr;
  for (size_t i = 0; i < n; ++i) {
    uint64_t d = 0;
    size_t bit = i * bw;
    for (uint32_t b = 0; b < bw; ++b, ++bit) d |= (uint64_t)((data[bit >> 3] >> (bit & 7)) & 1) << b;
    v[i] = (d + fr) & wmask(w);
  }
  *n_out = n;
  return 0;
}

/*
 * Decompress one partition whose element width is `w` bytes (the batch-wide
 * width, taken by the caller from partition 0's type byte like the reference).
 * Returns the per-partition status; *actual gets the decompressed byte count
 * (0 on failure).
 */
int casc_oracle_decompress(const void* comp_v, size_t comp_bytes, void* out_v, size_t out_cap, int w,
                           size_t* actual)
{
  const uint8_t* comp = (const uint8_t*)comp_v;
  uint8_t* out = (uint8_t*)out_v;
  *actual = 0;
  if (!comp || comp_bytes < 8 || (w != 1 && w != 2 && w != 4 && w != 8)) return ST_CANNOT_DECOMPRESS;
  const size_t end = comp_bytes / 4 * 4;
  const int R = comp[0], D = comp[1], bp = comp[2];
  uint32_t ub;
  memcpy(&ub, comp + 4, 4);
  const size_t n_total = ub / (size_t)w;
  if (out_cap < n_total * (size_t)w) return ST_CANNOT_DECOMPRESS;
  const size_t hdr = round_up(8, (size_t)w);
  if (R == 0 && D == 0 && bp == 0) {
    if (comp_bytes < hdr + n_total * (size_t)w) return ST_CANNOT_DECOMPRESS;
    memcpy(out, comp + hdr, n_total * (size_t)w);
    *actual = n_total * (size_t)w;
    return ST_SUCCESS;
  }
  if (R > 7) return ST_CANNOT_DECOMPRESS;
  const size_t meta_size = casc_chunk_meta_size(w, R, D);
  if (meta_size > 64) return ST_CANNOT_DECOMPRESS;
  const size_t chunk_elems = CASC_CHUNK_BYTES / (size_t)w;
  const size_t delta_off = round_up(4 + 4 * (size_t)(R + 1), (size_t)w);
  uint64_t a[CASC_CHUNK_BYTES], b[CASC_CHUNK_BYTES], cnt[CASC_CHUNK_BYTES];
  size_t done = 0, chunk = hdr;
  int ok = 1;
  while (chunk < end) {
    if (chunk + meta_size > end) { ok = 0; break; }
    uint32_t meta[9];
    memcpy(meta, comp + chunk, 4 * (size_t)(R + 2));
    const size_t chunk_bytes = meta[0];
    size_t off[9];
    off[0] = 0;
    for (int r = 0; r + 1 < R; ++r) off[r + 1] = round_up(off[r] + meta[r + 1], 4);
    if (R > 0) off[R] = round_up(off[R - 1] + meta[R], max_sz(4, (size_t)w));
    const size_t arrays = chunk + meta_size;
    size_t n;
    if (read_array(comp, arrays + off[R], meta[R + 1], end, w, bp, a, chunk_elems, &n) != 0) { ok = 0; break; }
    /* Undo the layers in the exact reverse of the encoder's order
     * (R1 D1 R2 D2 ...).  The reference interleaves by comparing the remaining
     * counts (:1332-1391), which is the same order whenever num_deltas <=
     * num_RLEs or num_RLEs == 0 and is not an inverse otherwise (it cannot
     * decode its own output for e.g. RLE=1, delta=2); see DESIGN.md. */
    const int layers = R > D ? R : D;
    uint64_t *cur = a, *oth = b;
    for (int l = layers - 1; l >= 0 && ok; --l) {
      if (l < D) {
        if (n + 1 > chunk_elems) { ok = 0; break; }
        uint64_t acc = ld_elem(comp + chunk + delta_off + (size_t)l * (size_t)w, w);
        for (size_t i = 0; i < n; ++i) { oth[i] = acc; acc = (acc + cur[i]) & wmask(w); }
        oth[n] = acc;
        uint64_t* t = cur; cur = oth; oth = t;
        ++n;
      }
      if (l < R) {
        size_t nc;
        if (read_array(comp, arrays + off[l], meta[l + 1], end, 2, bp, cnt, chunk_elems, &nc) != 0
            || nc < n) { ok = 0; break; }
        size_t o = 0;
        for (size_t i = 0; i < n && ok; ++i) {
          if (o + cnt[i] > chunk_elems) { ok = 0; break; }
          for (size_t k = 0; k < cnt[i]; ++k) oth[o++] = cur[i];
        }
        if (!ok) break;
        uint64_t* t = cur; cur = oth; oth = t;
        n = o;
      }
    }
    if (!ok) break;
    if (done + n > n_total) { ok = 0; break; }
    for (size_t i = 0; i < n; ++i) st_elem(out + (done + i) * (size_t)w, w, cur[i]);
    done += n;
    if (chunk_bytes < meta_size) { ok = 0; break; } /* would not advance */
    chunk = round_up(chunk + chunk_bytes / 4 * 4, (size_t)w);
  }
  if (done != n_total) ok = 0;
  if (!ok) return ST_CANNOT_DECOMPRESS;
  *actual = done * (size_t)w;
  return ST_SUCCESS;
}

size_t casc_oracle_get_size(const void* comp, size_t comp_bytes)
{
  if (comp_bytes < 8) return 0;
  uint32_t ub;
  memcpy(&ub, (const uint8_t*)comp + 4, 4);
  return ub;
}

/* batch drivers used by the timed CPU baseline (one call, many partitions) */
void casc_oracle_compress_batch(const uint8_t* in, const size_t* in_off, const size_t* in_bytes, size_t batch,
                                uint8_t* out, const size_t* out_off, size_t* out_bytes, int type, int R, int D,
                                int bp)
{
  for (size_t i = 0; i < batch; ++i)
    out_bytes[i] = casc_oracle_compress(in + in_off[i], in_bytes[i], out + out_off[i], type, R, D, bp);
}

void casc_oracle_decompress_batch(const uint8_t* comp, const size_t* comp_off, const size_t* comp_bytes,
                                  size_t batch, uint8_t* out, const size_t* out_off, const size_t* out_cap, int w,
                                  size_t* actual, int* status)
{
  for (size_t i = 0; i < batch; ++i)
    status[i] = casc_oracle_decompress(comp + comp_off[i], comp_bytes[i], out + out_off[i], out_cap[i], w,
                                       &actual[i]);
}
