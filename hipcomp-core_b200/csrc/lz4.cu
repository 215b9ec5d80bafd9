// b200comp — batched LZ4 block codec for sm_100a.
//
// C ABI replaces reference src/lowlevel/LZ4Batch.cpp:71-224.
//
// Encoders.  Chunks up to 64 KiB (every headline configuration) go to the team encoder of
// lz_team_encode.cuh: one CTA per chunk, chunk and hash table resident in shared memory, all
// positions matched in parallel.  Its streams are plain LZ4 blocks with a ratio at or above the
// reference's, but not the reference's bytes.  Larger chunks (up to the format's 16 MiB) keep
// the one-warp-per-chunk kernel below, which reproduces the reference's greedy policy
// (src/LZ4Kernels.hiph:793-969: 32-position window, intra-window twins, uint16 hash table in
// the caller's workspace, last-lane-wins insertion) byte for byte.
// The decoder bounds every compressed read (the reference's does not).
#include "lz_common.cuh"
#include "lz_team_encode.cuh"

#include <mutex>

#pragma GCC visibility push(default)
#include "hipcomp/lz4.h"
#include "hipcomp/b200_ext.h"
#pragma GCC visibility pop

namespace b200comp {
namespace lz4 {

constexpr uint32_t kNullOffset = 0xFFFFu;
constexpr uint32_t kMaxOffset = 65535u;
constexpr size_t kMaxChunk = 1u << 24;
constexpr uint32_t kMaxHashTable = 1u << 14;

static size_t hash_table_size(size_t max_chunk)
{
  size_t p = 1;
  while (p < max_chunk) p *= 2;
  return p < kMaxHashTable ? p : kMaxHashTable;
}

static size_t max_compressed(size_t n) { return round_up_sz(n + 1 + (n + 254) / 255, 8); }

__device__ __forceinline__ uint32_t hash_key(uint32_t key, uint32_t ht) { return (__brev(key) + (key ^ 0xc375u)) & (ht - 1); }

// length-in-sequence-of-0xFF encoding, written cooperatively
__device__ __forceinline__ uint32_t put_lsic(uint8_t* out, uint32_t x)
{
  const uint32_t num = x / 255u + 1u;
  for (uint32_t i = (uint32_t)lane_id(); i < num; i += 32) out[i] = (i + 1 < num) ? (uint8_t)0xff : (uint8_t)(x % 255u);
  return num;
}

// token, literal length, literals, offset, match length (LZ4Kernels.hiph:665-715)
__device__ __forceinline__ uint32_t put_sequence(uint8_t* out, uint32_t c, const uint8_t* lit, uint32_t nlit,
                                                 uint32_t nmatch, uint32_t offset)
{
  if (lane_id() == 0) {
    const uint32_t hi = nlit >= 15 ? 15u : nlit;
    const uint32_t lo = nmatch >= 19 ? 15u : ((nmatch - 4u) & 0x0fu); // (0-4)&15 on the last token, as the reference
    out[c] = (uint8_t)((hi << 4) | lo);
  }
  ++c;
  if (nlit >= 15) c += put_lsic(out + c, nlit - 15);
  warp_copy_ro(out + c, lit, nlit);
  c += nlit;
  if (nmatch > 0) {
    if (lane_id() == 0) {
      out[c] = (uint8_t)(offset & 0xff);
      out[c + 1] = (uint8_t)(offset >> 8);
    }
    c += 2;
    if (nmatch >= 19) c += put_lsic(out + c, nmatch - 19);
  }
  return c;
}

constexpr int kCompWarps = 4;

// One warp per chunk, kCompWarps chunks per CTA.  TS = match granularity in bytes
// (data_type width).  kGlobalTable: the uint16 hash table of a chunk lives in the
// caller's workspace (L2-resident, as many warps per SM as registers allow) instead
// of shared memory (chosen for tables <= 8 KiB, where 28 warps/SM still fit).
//
// Per window the dependent chain is: keys (1 load round) -> table probe, issued for
// every valid lane before the intra-window twin search finishes -> candidate verify
// -> word-wide match extension.  Literals that lie inside the current window are
// stored straight from the key registers.
#ifndef B200_LZ4_COMP_MINB
#define B200_LZ4_COMP_MINB 10
#endif
template <int TS, bool kGlobalTable>
__global__ void __launch_bounds__(kCompWarps * 32, B200_LZ4_COMP_MINB)
compress_kernel(const void* const* in_ptrs, const size_t* in_bytes, void* const* out_ptrs, size_t* out_bytes,
                size_t batch, uint32_t ht_size, uint16_t* gtables)
{
  extern __shared__ __align__(16) uint16_t stables[];
  const uint32_t lane = (uint32_t)lane_id();
  const int warp = threadIdx.x >> 5;
  constexpr uint32_t last_valid_match = (12 + TS - 1) / TS;
  constexpr uint32_t min_ending = (5 + TS - 1) / TS;
  constexpr uint32_t window = 32 - 3 / TS;
  const size_t stride = (size_t)gridDim.x * kCompWarps;

  for (size_t chunk = (size_t)blockIdx.x * kCompWarps + warp; chunk < batch; chunk += stride) {
    const uint8_t* in = static_cast<const uint8_t*>(in_ptrs[chunk]);
    uint8_t* out = static_cast<uint8_t*>(out_ptrs[chunk]);
    const uint32_t length = (uint32_t)in_bytes[chunk];
    const uint32_t typed_length = (length + TS - 1) / TS;
    // global tables belong to the (persistent) warp, not to the chunk: the set of live tables is
    // gridDim.x * kCompWarps * 2 * ht_size bytes, sized by the launcher to stay inside the L2
    uint16_t* table = kGlobalTable ? gtables + ((size_t)blockIdx.x * kCompWarps + warp) * ht_size
                                   : stables + (size_t)warp * ht_size;

    if (ht_size >= 8 && ((reinterpret_cast<uintptr_t>(table) & 15) == 0)) {
      uint4* t128 = reinterpret_cast<uint4*>(table);
      const uint4 ones = make_uint4(~0u, ~0u, ~0u, ~0u);
      for (uint32_t i = lane; i < ht_size / 8; i += 32) t128[i] = ones;
    } else {
      for (uint32_t i = lane; i < ht_size; i += 32) table[i] = (uint16_t)kNullOffset;
    }
    __syncwarp();

    uint32_t d = 0, c = 0;
    while (d < typed_length) {
      const uint32_t token_start = d;
      for (;;) {
        if (d + last_valid_match >= typed_length) {
          c = put_sequence(out, c, in + (size_t)token_start * TS, length - token_start * TS, 0, 0);
          d = typed_length;
          break;
        }
        const uint32_t nvalid = min(window, typed_length - d - last_valid_match);
        const unsigned vmask = nvalid >= 32 ? kFull : ((1u << nvalid) - 1u);
        const bool valid = lane < nvalid;
        const uint32_t key = valid ? ld_unaligned32_ro(in + (size_t)(d + lane) * TS) : 0u;
        const uint32_t hpos = hash_key(key, ht_size);
        // speculative probe: does not depend on the twin search below
        const uint32_t h = valid ? (uint32_t)table[hpos] : kNullOffset;

        // (1) earliest lane that repeats the key of a lower lane in this window
        unsigned same = 0;
        if (valid) same = __match_any_sync(vmask, key);
        const uint32_t lowest = same ? (uint32_t)__ffs((int)same) - 1u : lane;
        const unsigned twins = __ballot_sync(kFull, valid && lowest != lane);
        uint32_t first = nvalid;
        uint32_t match_loc = typed_length;
        if (twins) {
          first = (uint32_t)__ffs((int)twins) - 1u;
          match_loc = d + __shfl_sync(kFull, lowest, (int)first);
        }
        // (2) lanes in front of it verify their table candidate
        bool found = false;
        uint32_t cand = 0;
        if (lane < first && h != kNullOffset) {
          const uint32_t pos = d + lane;
          cand = (pos & ~kMaxOffset) + h;
          if (cand >= pos) cand -= 65536u;
          const uint32_t dist = pos - cand;
          if (cand < pos && dist <= kMaxOffset && dist * TS <= kMaxOffset)
            found = ld_unaligned32_ro(in + (size_t)cand * TS) == key;
        }
        const unsigned hits = __ballot_sync(kFull, found);
        if (hits) {
          first = (uint32_t)__ffs((int)hits) - 1u;
          match_loc = __shfl_sync(kFull, cand, (int)first);
        }
        // (3) insert the lanes in front of the match (all valid lanes when there
        //     is none); the highest lane wins a contended slot
        const uint32_t ninsert = match_loc != typed_length ? first : nvalid;
        if (lane < ninsert) {
          const unsigned imask = ninsert >= 32 ? kFull : ((1u << ninsert) - 1u);
          const unsigned grp = __match_any_sync(imask, hpos);
          if (31u - (uint32_t)__clz((int)grp) == lane) table[hpos] = (uint16_t)((d + lane) & kMaxOffset);
        }
        __syncwarp();
        if (match_loc == typed_length) {
          d += nvalid;
          continue;
        }
        // (4) extend the match: 4 bytes per lane, 128 bytes per step
        const uint32_t pos = d + first;
        const uint32_t limit_b = (typed_length - pos - min_ending) * TS;
        const uint8_t* pa = in + (size_t)match_loc * TS;
        const uint8_t* pb = in + (size_t)pos * TS;
        uint32_t mlen_b = limit_b;
        for (uint32_t j = 0; j < limit_b; j += 128) {
          const uint32_t o = j + 4 * lane;
          uint32_t x = 1; // lanes past the limit: mismatch at their first byte
          if (o < limit_b) {
            x = ld_unaligned32_ro(pa + o) ^ ld_unaligned32_ro(pb + o);
            const uint32_t nb = limit_b - o;
            if (nb < 4) x |= 1u << (8 * nb);
          }
          const unsigned diff = __ballot_sync(kFull, x != 0);
          if (diff) {
            const uint32_t l = (uint32_t)__ffs((int)diff) - 1u;
            const uint32_t xb = __shfl_sync(kFull, x, (int)l);
            mlen_b = j + 4 * l + (((uint32_t)__ffs((int)xb) - 1u) >> 3);
            break;
          }
        }
        const uint32_t mlen = mlen_b / TS;
        const uint32_t nlit = pos - token_start;
        if (token_start == d) {
          // every literal of this sequence is a lane of the current window
          const uint32_t nlit_b = nlit * TS, nmatch_b = mlen * TS, off_b = (pos - match_loc) * TS;
          const uint32_t lit_at = c + 1 + (nlit_b >= 15 ? 1u : 0u); // nlit_b <= 116 < 15 + 255
          if (lane < first) {
#pragma unroll
            for (int b = 0; b < TS; ++b) out[lit_at + lane * TS + b] = (uint8_t)(key >> (8 * b));
          }
          if (lane == 0) {
            const uint32_t hi = nlit_b >= 15 ? 15u : nlit_b;
            const uint32_t lo = nmatch_b >= 19 ? 15u : (nmatch_b - 4u);
            out[c] = (uint8_t)((hi << 4) | lo);
            if (nlit_b >= 15) out[c + 1] = (uint8_t)(nlit_b - 15);
            out[lit_at + nlit_b] = (uint8_t)(off_b & 0xff);
            out[lit_at + nlit_b + 1] = (uint8_t)(off_b >> 8);
          }
          c = lit_at + nlit_b + 2;
          if (nmatch_b >= 19) c += put_lsic(out + c, nmatch_b - 19);
        } else {
          c = put_sequence(out, c, in + (size_t)token_start * TS, nlit * TS, mlen * TS, (pos - match_loc) * TS);
        }
        d = token_start + nlit + mlen;
        break;
      }
    }
    if (lane == 0) out_bytes[chunk] = c;
    __syncwarp();
  }
}

// ------------------------------------------------------------------ decode

constexpr int kDecWarps = 4;

// length field in 0xFF bytes, read from global memory with a bound on the stream and on the value
__device__ __forceinline__ bool read_lsic_bounded(const uint8_t* comp, uint32_t comp_len, uint32_t& ci,
                                                  uint32_t& value)
{
  uint32_t b = 0xff;
  while (b == 0xff) {
    if (ci >= comp_len || value > 0x7FFFFFFFu) return false;
    b = __ldg(comp + ci);
    ++ci;
    value += b;
  }
  return true;
}

// The decoder: one warp per chunk, rounds of up to 32 sequences.  The serial token chain is
// reduced to its minimum: a round stages 512 compressed bytes in shared memory (one
// 16-byte load per lane) and, from the same registers, computes for EVERY byte position
// the distance to the next token if a token started there (4 positions per SIMD word
// operation).  The serial part is then "idx += step[idx]", one shared-memory byte load
// per sequence.  Tokens with extended lengths (a 0xF nibble), the last token of the
// stream and tokens that cross the window end take a generic path that reads the same
// window; a sequence whose header does not fit a fresh window (literal runs of hundreds
// of bytes) is run straight from global memory with vector copies.  After the walk lane
// k decodes sequence k on its own, and the round executor (lz_common.cuh) produces the
// output 64 bytes per iteration.
constexpr int kWin = 512;
constexpr uint32_t kWinKeep = 224; // a staged window is reused while at least this much of it lies ahead

struct SlowRec {
  uint32_t lit_off, nlit, mlen, off;
};

// kSizeOnly: the same walk without the output (hipcompBatchedLZ4GetDecompressSizeAsync): `actual` receives the
// decompressed size of every chunk, 0 for a stream that does not parse.
template <bool kSizeOnly>
__global__ void __launch_bounds__(kDecWarps * 32)
decompress_window_kernel(const void* const* comp_ptrs, const size_t* comp_bytes, const size_t* out_caps,
                         void* const* out_ptrs, size_t* actual, hipcompStatus_t* statuses, size_t batch)
{
  __shared__ __align__(16) uint8_t s_win[kDecWarps][kWin + 32]; // walk may peek <= 18 bytes past the window
  __shared__ __align__(16) uint8_t s_step[kDecWarps][kWin];
  __shared__ SlowRec s_rec[kDecWarps][32];
  __shared__ uint32_t s_own[kDecWarps][kOwnWords];           // start bitmap, bases and stage of the round executor
  __shared__ uint16_t s_idx[kDecWarps][32];                  // token positions found by the fast walk
  const uint32_t lane = (uint32_t)lane_id();
  const int warp = threadIdx.x >> 5;
  uint8_t* win = s_win[warp];
  uint8_t* step = s_step[warp];
  SlowRec* rec = s_rec[warp];
  uint16_t* idxs = s_idx[warp];
  const size_t stride = (size_t)gridDim.x * kDecWarps;
  for (size_t chunk = (size_t)blockIdx.x * kDecWarps + warp; chunk < batch; chunk += stride) {
    const uint8_t* comp = static_cast<const uint8_t*>(comp_ptrs[chunk]);
    uint8_t* out = kSizeOnly ? nullptr : static_cast<uint8_t*>(out_ptrs[chunk]);
    const uint32_t comp_len = (uint32_t)comp_bytes[chunk];
    const size_t cap64 = kSizeOnly ? 0xFFFFFFFEull : out_caps[chunk]; // size query: anything below 4 GiB
    const uint32_t cap = cap64 > 0xFFFFFFFFull ? 0xFFFFFFFFu : (uint32_t)cap64;
    const uintptr_t cbase = reinterpret_cast<uintptr_t>(comp);
    uint32_t ci = 0, di = 0;
    bool ok = true;
    // the staged window [wpos, wpos + wvalid) and its step table stay valid across rounds: a round of 32
    // sequences consumes ~200 of the 512 bytes, so about every second round starts without a load from global memory
    uint32_t wpos = 0, wvalid = 0;
    bool staged = false;
    while (ok && ci < comp_len) {
      uint32_t shift = ci - wpos; // where the walk starts inside the window (wpos may lie up to 15 bytes below 0: wrap-safe)
      const bool reuse = staged && shift < wvalid && (shift + kWinKeep <= wvalid || wpos + wvalid == comp_len);
      const bool fresh = !reuse;
      if (fresh) {
      // ---- stage [ci - shift, ci - shift + 512): 16-byte aligned blocks that intersect the stream
      shift = (uint32_t)((cbase + ci) & 15);
      wpos = ci - shift; // stream position of window byte 0 (may "underflow" by < 16 at ci = 0)
      wvalid = min((uint32_t)kWin, comp_len - wpos);
      staged = true;
      __syncwarp(); // the previous round's readers of the window are done
      uint4 v = make_uint4(0, 0, 0, 0);
      if (16u * lane < wvalid) v = ld_block16_guarded(cbase, comp_len, cbase + ci - shift + 16u * lane, win + 16 * lane);
      *reinterpret_cast<uint4*>(win + 16 * lane) = v;
      {
        // step[p], for a token at p: 3 + literal nibble (token, literals, offset); bit 7 set when the
        // match nibble is 0xF (a length byte follows the offset); 0 when the literal nibble is 0xF
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint32_t r[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t nl = (w[j] >> 4) & 0x0f0f0f0fu, mc = w[j] & 0x0f0f0f0fu;
          const uint32_t nl15 = ((nl + 0x01010101u) >> 4) & 0x01010101u;
          const uint32_t mc15 = ((mc + 0x01010101u) << 3) & 0x80808080u;
          r[j] = ((nl + 0x03030303u) | mc15) & ~(nl15 * 0xffu);
        }
        *reinterpret_cast<uint4*>(step + 16 * lane) = make_uint4(r[0], r[1], r[2], r[3]);
      }
      } // fresh
      __syncwarp();
      const bool at_stream_end = wpos + wvalid == comp_len;

      // ---- walk the token chain.  Fast part: tokens that start at least 19 bytes in front of the window's end
      // are whole sequences inside the window (a step is at most 18), one table byte each.
      uint32_t idx = shift, k = 0, my_idx = 0;
      bool my_slow = false, giant = false;
      {
        const uint32_t wfast = wvalid >= 20u ? wvalid - 19u : 0u;
#pragma unroll 8
        for (; k < 32u; ++k) {
          if (idx >= wfast) break;
          const uint32_t s = step[idx];
          uint32_t adv = s & 0x7fu;
          if (s & 0x80u) {
            // match length 19..273: exactly one length byte unless it reads 0xFF
            const uint32_t e = win[idx + adv];
            adv = e == 0xffu ? 0u : adv + 1u;
          }
          if (adv == 0u) break;
          idxs[k] = (uint16_t)idx;
          idx += adv;
        }
      }
      const uint32_t kfast = k;
      // the rest of the round: tokens near the window's end, extended lengths, the last sequence
      while (k < 32u && idx < wvalid) {
        const uint32_t s = step[idx];
        uint32_t adv = s & 0x7fu;
        if (s & 0x80u) {
          const uint32_t e = win[idx + adv];
          adv = e == 0xffu ? 0u : adv + 1u;
        }
        if (adv != 0u && idx + adv <= wvalid) {
          if (lane == k) my_idx = idx;
          idx += adv;
          ++k;
          continue;
        }
        // generic parse of the sequence at idx, from the window
        bool more = false; // header or literals reach past the window
        uint32_t p = idx;
        const uint32_t tok = win[p++];
        uint32_t nlit = tok >> 4, mlen = 0, off = 0;
        if (nlit == 15u) {
          uint32_t b = 0xff;
          while (b == 0xffu) {
            if (p >= wvalid) { more = true; break; }
            if (nlit > 0x7FFFFFFFu) { ok = false; break; }
            b = win[p++];
            nlit += b;
          }
        }
        if (!ok) break;
        uint32_t lit_off = p;
        if (!more) {
          const uint32_t left = comp_len - (wpos + p); // stream bytes from the first literal on
          if (nlit > left) { ok = false; break; }
          if (nlit == left) {
            // last sequence of the stream: literals only
            if (p + nlit > wvalid) more = true;
            else p += nlit;
          } else {
            if (nlit + 2u > left) { ok = false; break; }
            if (p + nlit + 2u > wvalid) more = true;
            else {
              off = (uint32_t)win[p + nlit] | ((uint32_t)win[p + nlit + 1] << 8);
              p += nlit + 2u;
              mlen = 4u + (tok & 15u);
              if ((tok & 15u) == 15u) {
                uint32_t b = 0xff;
                while (b == 0xffu) {
                  if (p >= wvalid) { more = true; break; }
                  if (mlen > 0x7FFFFFFFu) { ok = false; break; }
                  b = win[p++];
                  mlen += b;
                }
              }
              if (!ok) break;
              if (!more && off == 0u) { ok = false; break; }
            }
          }
        }
        if (more) {
          if (at_stream_end) {
            // the window already holds the rest of the stream: the header is truncated
            // (literal runs that pass the end were diagnosed above)
            if (!(k == 0u && fresh)) break; // re-stage at this token first: cheap and uniform
            ok = false;
            break;
          }
          if (k == 0u && fresh) giant = true;
          break;
        }
        if (lane == k) {
          my_idx = idx;
          my_slow = true;
          rec[k] = SlowRec{lit_off, nlit, mlen, off};
        }
        idx = p;
        ++k;
      }
      if (!ok) break;
      if (k == 0u && !fresh) { // the token did not fit what is left of a reused window: stage at it
        staged = false;
        continue;
      }
      if (giant) {
        // one sequence straight from global memory (long literal run and / or long headers)
        const uint32_t tok = __ldg(comp + ci);
        ++ci;
        uint32_t nlit = tok >> 4;
        if (nlit == 15 && !read_lsic_bounded(comp, comp_len, ci, nlit)) { ok = false; break; }
        if (nlit > comp_len - ci || nlit > cap - di) { ok = false; break; }
        if (!kSizeOnly) group_copy_ro<32>(out + di, comp + ci, nlit, lane);
        ci += nlit;
        di += nlit;
        if (ci < comp_len) {
          if (ci + 2 > comp_len) { ok = false; break; }
          const uint32_t off = (uint32_t)__ldg(comp + ci) | ((uint32_t)__ldg(comp + ci + 1) << 8);
          ci += 2;
          uint32_t mlen = 4 + (tok & 15);
          if ((tok & 15) == 15 && !read_lsic_bounded(comp, comp_len, ci, mlen)) { ok = false; break; }
          if (off == 0 || off > di || mlen > cap - di) { ok = false; break; }
          __syncwarp();
          if (!kSizeOnly) group_copy_match<32>(out + di, off, mlen, lane, kFull);
          di += mlen;
        }
        __syncwarp();
        continue;
      }
      if (k == 0u) {
        // at_stream_end with a truncated header after a re-stage lands here as k == 0 && idx == shift
        ok = false;
        break;
      }
      // ---- lane j decodes sequence j
      __syncwarp();
      if (lane < kfast) my_idx = idxs[lane];
      uint32_t my_nlit = 0, my_mlen = 0, my_off = 0, my_lp = 0;
      if (lane < k) {
        if (my_slow) {
          const SlowRec r = rec[lane];
          my_nlit = r.nlit;
          my_mlen = r.mlen;
          my_off = r.off;
          my_lp = wpos + r.lit_off;
        } else {
          const uint32_t tok = win[my_idx];
          my_nlit = tok >> 4;
          my_mlen = 4u + (tok & 15u);
          const uint32_t op = my_idx + 1u + my_nlit;
          my_off = (uint32_t)win[op] | ((uint32_t)win[op + 1u] << 8);
          if ((tok & 15u) == 15u) my_mlen += win[op + 2u];
          my_lp = wpos + my_idx + 1u;
        }
      }
      const bool zero_off = lane < k && my_mlen != 0u && my_off == 0u;
      uint32_t total;
      const uint32_t pos = add_sat_u32(di, warp_excl_sum_sat(add_sat_u32(my_nlit, my_mlen), total));
      const uint32_t lit_end = add_sat_u32(pos, my_nlit);
      const uint32_t seq_end = add_sat_u32(lit_end, my_mlen);
      const bool bad = zero_off || seq_end > cap || (my_mlen != 0 && my_off > lit_end);
      if (__any_sync(kFull, bad)) { ok = false; break; }
      const uint32_t end = di + total;
      // literal byte p of a sequence is comp[p + (my_lp - pos)], i.e. window byte p + (my_lp - wpos - pos)
      if (!kSizeOnly) {
        if (!execute_round_owned(out, win, s_own[warp], di, end, pos, lit_end, seq_end, my_lp - wpos - pos, my_off))
          execute_round(out, comp, di, end, pos, lit_end, seq_end, my_lp - pos, my_off);
      }
      di = end;
      ci = wpos + idx;
      __syncwarp(); // the window is rewritten next
    }
    if (lane == 0) {
      if (actual) actual[chunk] = ok ? di : 0;
      if (!kSizeOnly && statuses) statuses[chunk] = ok ? hipcompSuccess : hipcompErrorCannotDecompress;
    }
    __syncwarp();
  }
}

// ---- team encoder (lz_team_encode.cuh): how an LZ4 sequence is sized and written.  Hash-table bits and team
//      width by declared chunk size; oracle/lz_team_model.c is called with the same hash_log by the tests.
constexpr size_t kTeamMaxChunk = 65536;

inline size_t compress_temp_bytes(size_t max_chunk, size_t batch)
{
  if (max_chunk <= kTeamMaxChunk) return batch ? 8 : 0;
  return hash_table_size(max_chunk) * sizeof(uint16_t) * batch;
}

struct TeamEmit {
  static constexpr bool kNeedDist = false;
  static __device__ __forceinline__ uint32_t lsic_bytes(uint32_t x) { return x >= 15u ? (x - 15u) / 255u + 1u : 0u; }
  static __device__ __forceinline__ uint32_t empty_stream(uint8_t*) { return 0; }
  static __device__ __forceinline__ uint32_t preamble(uint8_t*, uint32_t, uint32_t) { return 0; }
  static __device__ __forceinline__ uint32_t lit_bytes(uint32_t nlit) { return nlit + lsic_bytes(nlit); }
  // token + offset + match length bytes
  static __device__ __forceinline__ uint32_t match_bytes(uint32_t len, uint32_t) { return 3u + lsic_bytes(len - 4u); }
  static __device__ __forceinline__ uint8_t* put_lit_header(uint8_t* o, uint32_t nlit, uint32_t len)
  {
    const uint32_t ml = len - 4u;
    *o++ = (uint8_t)(((nlit >= 15u ? 15u : nlit) << 4) | (ml >= 15u ? 15u : ml));
    if (nlit >= 15u) {
      uint32_t x = nlit - 15u;
      for (; x >= 255u; x -= 255u) *o++ = 0xff;
      *o++ = (uint8_t)x;
    }
    return o;
  }
  static __device__ __forceinline__ uint8_t* put_match(uint8_t* o, uint32_t d, uint32_t len)
  {
    o[0] = (uint8_t)(d & 0xffu);
    o[1] = (uint8_t)(d >> 8);
    o += 2;
    const uint32_t ml = len - 4u;
    if (ml >= 15u) {
      uint32_t x = ml - 15u;
      for (; x >= 255u; x -= 255u) *o++ = 0xff;
      *o++ = (uint8_t)x;
    }
    return o;
  }
  // the match part of a long match (len > team::kLongMatch), written by the whole warp
  static __device__ __forceinline__ void put_match_warp(uint8_t* o, uint32_t d, uint32_t len, uint32_t lane)
  {
    if (lane == 0) {
      o[0] = (uint8_t)(d & 0xffu);
      o[1] = (uint8_t)(d >> 8);
    }
    team::put_lsic(o + 2, len - 4u - 15u);
  }
  // header of the final literals, written by the whole warp; returns its size
  static __device__ __forceinline__ uint32_t put_tail_header(uint8_t* o, uint32_t nlit, uint32_t lane)
  {
    if (lane == 0) o[0] = (uint8_t)((nlit >= 15u ? 15u : nlit) << 4);
    return 1u + (nlit >= 15u ? team::put_lsic(o + 1, nlit - 15u) : 0u);
  }
};

template <int CAP, int HL, int W>
static void launch_team_cfg(const void* const* in_ptrs, const size_t* in_bytes, size_t batch, void* const* out_ptrs,
                            size_t* out_bytes, cudaStream_t stream, int ctas_per_sm)
{
  using L = team::TeamLayout<CAP, HL, W>;
  allow_smem(team::compress_team_kernel<TeamEmit, CAP, HL, W>, (int)L::kBytes);
  const size_t cap = (size_t)sm_count() * ctas_per_sm;
  const int grid = (int)(batch < cap ? batch : cap);
  team::compress_team_kernel<TeamEmit, CAP, HL, W><<<grid, W * 32, L::kBytes, stream>>>(in_ptrs, in_bytes, out_ptrs, out_bytes, batch);
}

static void launch_team(const void* const* in_ptrs, const size_t* in_bytes, size_t max_chunk, size_t batch,
                        void* const* out_ptrs, size_t* out_bytes, cudaStream_t stream)
{
  if (max_chunk <= 4096)
    launch_team_cfg<4096, 12, 4>(in_ptrs, in_bytes, batch, out_ptrs, out_bytes, stream, 6);
  else if (max_chunk <= 16384)
    launch_team_cfg<16384, 13, 8>(in_ptrs, in_bytes, batch, out_ptrs, out_bytes, stream, 3);
  else
    launch_team_cfg<65536, 14, 24>(in_ptrs, in_bytes, batch, out_ptrs, out_bytes, stream, 1);
}

} // namespace lz4
} // namespace b200comp

using namespace b200comp;

extern "C" {

hipcompStatus_t hipcompBatchedLZ4CompressGetTempSize(size_t batch_size, size_t max_uncompressed_chunk_bytes,
                                                     hipcompBatchedLZ4Opts_t, size_t* temp_bytes)
{
  static const char* fn = "hipcompBatchedLZ4CompressGetTempSize()";
  B200_CHECK_NOT_NULL(temp_bytes, fn);
  if (max_uncompressed_chunk_bytes > lz4::kMaxChunk)
    return fail(hipcompErrorInvalidValue, fn, "Maximum chunk size for LZ4 is " + std::to_string(lz4::kMaxChunk));
  // Chunks up to 64 KiB keep their hash table in shared memory (team encoder): no workspace.  Above, the
  // reference's figure (a uint16 table per chunk, LZ4CompressionKernels.hip:142-156).
  *temp_bytes = lz4::compress_temp_bytes(max_uncompressed_chunk_bytes, batch_size);
  return hipcompSuccess;
}

hipcompStatus_t hipcompBatchedLZ4CompressGetMaxOutputChunkSize(size_t max_uncompressed_chunk_bytes,
                                                               hipcompBatchedLZ4Opts_t, size_t* max_compressed_bytes)
{
  static const char* fn = "hipcompBatchedLZ4CompressGetOutputSize()";
  B200_CHECK_NOT_NULL(max_compressed_bytes, fn);
  if (max_uncompressed_chunk_bytes > lz4::kMaxChunk)
    return fail(hipcompErrorInvalidValue, fn, "Maximum chunk size for LZ4 is " + std::to_string(lz4::kMaxChunk));
  *max_compressed_bytes = lz4::max_compressed(max_uncompressed_chunk_bytes);
  return hipcompSuccess;
}

hipcompStatus_t hipcompBatchedLZ4CompressAsync(const void* const* device_uncompressed_ptrs,
                                               const size_t* device_uncompressed_bytes,
                                               size_t max_uncompressed_chunk_bytes, size_t batch_size,
                                               void* device_temp_ptr, size_t temp_bytes,
                                               void* const* device_compressed_ptrs, size_t* device_compressed_bytes,
                                               hipcompBatchedLZ4Opts_t format_opts, hipStream_t stream)
{
  static const char* fn = "hipcompBatchedLZ4CompressAsync()";
  B200_TRACE(fn);
  if (max_uncompressed_chunk_bytes > lz4::kMaxChunk)
    return fail(hipcompErrorInvalidValue, fn, "Maximum chunk size for LZ4 is " + std::to_string(lz4::kMaxChunk));
  const uint32_t ht = (uint32_t)lz4::hash_table_size(max_uncompressed_chunk_bytes);
  const size_t required = lz4::compress_temp_bytes(max_uncompressed_chunk_bytes, batch_size);
  if (temp_bytes < required)
    return fail(hipcompErrorInvalidValue, fn,
                "Insufficient temp space: got " + std::to_string(temp_bytes) + " bytes, but need "
                    + std::to_string(required) + " bytes.");
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_NOT_NULL(device_uncompressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_compressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_ptrs, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_ptrs, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_bytes, fn);
  switch (format_opts.data_type) {
  case HIPCOMP_TYPE_BITS:
  case HIPCOMP_TYPE_CHAR:
  case HIPCOMP_TYPE_UCHAR:
  case HIPCOMP_TYPE_SHORT:
  case HIPCOMP_TYPE_USHORT:
  case HIPCOMP_TYPE_INT:
  case HIPCOMP_TYPE_UINT:
    break;
  default:
    return fail(hipcompErrorInvalidValue, fn, "Unsupported input data type");
  }
  if (max_uncompressed_chunk_bytes <= lz4::kTeamMaxChunk) {
    // byte-granular matching for every data type: a superset of the element-granular matches
    lz4::launch_team(device_uncompressed_ptrs, device_uncompressed_bytes, max_uncompressed_chunk_bytes,
                     batch_size, device_compressed_ptrs, device_compressed_bytes, stream);
    return check_launch(fn);
  }
  const bool global_table = (size_t)ht * sizeof(uint16_t) > 8192;
  if (global_table && device_temp_ptr == nullptr)
    return fail(hipcompErrorInvalidValue, fn, "'device_temp_ptr' must not be null.");
  const int smem = global_table ? 0 : (int)(ht * sizeof(uint16_t) * lz4::kCompWarps);
  const size_t need = (batch_size + lz4::kCompWarps - 1) / lz4::kCompWarps;
  // CTAs per SM: with global tables, few enough that every live table stays L2-resident
  // (the workspace always has room: grid * kCompWarps <= batch)
  const int ctas_per_sm = global_table ? B200_LZ4_COMP_MINB : 16;
  const size_t cap = (size_t)sm_count() * ctas_per_sm;
  const int grid = (int)(need < cap ? need : cap);
  uint16_t* gt = static_cast<uint16_t*>(device_temp_ptr);
#define B200_LZ4_LAUNCH(TS)                                                                                        \
  do {                                                                                                             \
    if (global_table)                                                                                              \
      lz4::compress_kernel<TS, true><<<grid, lz4::kCompWarps * 32, 0, stream>>>(                                   \
          device_uncompressed_ptrs, device_uncompressed_bytes, device_compressed_ptrs, device_compressed_bytes,    \
          batch_size, ht, gt);                                                                                     \
    else                                                                                                           \
      lz4::compress_kernel<TS, false><<<grid, lz4::kCompWarps * 32, smem, stream>>>(                               \
          device_uncompressed_ptrs, device_uncompressed_bytes, device_compressed_ptrs, device_compressed_bytes,    \
          batch_size, ht, gt);                                                                                     \
  } while (0)
  switch (format_opts.data_type) {
  case HIPCOMP_TYPE_BITS:
  case HIPCOMP_TYPE_CHAR:
  case HIPCOMP_TYPE_UCHAR:
    B200_LZ4_LAUNCH(1);
    break;
  case HIPCOMP_TYPE_SHORT:
  case HIPCOMP_TYPE_USHORT:
    B200_LZ4_LAUNCH(2);
    break;
  case HIPCOMP_TYPE_INT:
  case HIPCOMP_TYPE_UINT:
    B200_LZ4_LAUNCH(4);
    break;
  default:
    return fail(hipcompErrorInvalidValue, fn, "Unsupported input data type");
  }
#undef B200_LZ4_LAUNCH
  return check_launch(fn);
}

hipcompStatus_t hipcompBatchedLZ4DecompressGetTempSize(size_t num_chunks, size_t, size_t* temp_bytes)
{
  B200_CHECK_NOT_NULL(temp_bytes, "hipcompBatchedLZ4DecompressGetTempSize()");
  *temp_bytes = round_up_sz(24 * num_chunks, 8);
  return hipcompSuccess;
}

hipcompStatus_t hipcompBatchedLZ4DecompressAsync(const void* const* device_compressed_ptrs,
                                                 const size_t* device_compressed_bytes,
                                                 const size_t* device_uncompressed_bytes,
                                                 size_t* device_actual_uncompressed_bytes, size_t batch_size,
                                                 void* const, size_t, void* const* device_uncompressed_ptrs,
                                                 hipcompStatus_t* device_statuses, hipStream_t stream)
{
  static const char* fn = "hipcompBatchedLZ4DecompressAsync()";
  B200_TRACE(fn);
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_NOT_NULL(device_compressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_ptrs, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_ptrs, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_ptrs, fn);
  const size_t need = (batch_size + lz4::kDecWarps - 1) / lz4::kDecWarps;
  const size_t cap = (size_t)sm_count() * 16;
  lz4::decompress_window_kernel<false><<<(int)(need < cap ? need : cap), lz4::kDecWarps * 32, 0, stream>>>(
      device_compressed_ptrs, device_compressed_bytes, device_uncompressed_bytes, device_uncompressed_ptrs,
      device_actual_uncompressed_bytes, device_statuses, batch_size);
  return check_launch(fn);
}

hipcompStatus_t hipcompBatchedLZ4GetDecompressSizeAsync(const void* const* device_compressed_ptrs,
                                                        const size_t* device_compressed_bytes,
                                                        size_t* device_uncompressed_bytes, size_t batch_size,
                                                        hipStream_t stream)
{
  static const char* fn = "hipcompBatchedLZ4GetDecompressSizeAsync()";
  B200_TRACE(fn);
  B200_CHECK_NOT_NULL(device_compressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  if (batch_size == 0) return hipcompSuccess;
  // one warp per chunk walks the token chain through the decoder's staged window (a thread per chunk diverges
  // on every byte: C4 has 4 Mi chunks)
  const size_t need = (batch_size + lz4::kDecWarps - 1) / lz4::kDecWarps;
  const size_t cap = (size_t)sm_count() * 16;
  lz4::decompress_window_kernel<true><<<(int)(need < cap ? need : cap), lz4::kDecWarps * 32, 0, stream>>>(
      device_compressed_ptrs, device_compressed_bytes, nullptr, nullptr, device_uncompressed_bytes, nullptr, batch_size);
  return check_launch(fn);
}

} // extern "C"

// ---- Allocation-free compress entry point (include/hipcomp/b200_ext.h; no counterpart in the reference).
// The hash tables that hipcompBatchedLZ4CompressAsync keeps in the caller's workspace for chunks above 64 KiB
// live here in a static device array instead: one copy per GPU that loads the library, held until exit, shared
// by every call.  The encoder resets a chunk's table before it uses it, so nothing has to be cleared between
// calls; two calls that overlap on the GPU would share tables and race.
static __device__ __align__(256) unsigned char g_lz4_noalloc_scratch[B200COMP_LZ4_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[kNoAllocMaxDevices] = {};

// Device address of the scratch array on the current device.  The array's name is only a host-side stand-in in
// host code: the address a kernel can use comes from cudaGetSymbolAddress, taken once per device with that
// device current.
cudaError_t noalloc_scratch(void** out)
{
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kNoAllocMaxDevices) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_lz4_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return cudaSuccess;
}
} // namespace

extern "C" hipcompStatus_t b200compBatchedLZ4CompressNoAllocAsync(
    const void* const* device_uncompressed_ptrs, const size_t* device_uncompressed_bytes,
    size_t max_uncompressed_chunk_bytes, size_t batch_size, void* const* device_compressed_ptrs,
    size_t* device_compressed_bytes, hipcompBatchedLZ4Opts_t format_opts, hipStream_t stream)
{
  static const char* fn = "b200compBatchedLZ4CompressNoAllocAsync()";
  if (max_uncompressed_chunk_bytes > lz4::kMaxChunk)
    return fail(hipcompErrorInvalidValue, fn, "Maximum chunk size for LZ4 is " + std::to_string(lz4::kMaxChunk));
  const size_t required = lz4::compress_temp_bytes(max_uncompressed_chunk_bytes, batch_size);
  if (required > (size_t)B200COMP_LZ4_NOALLOC_SCRATCH_BYTES)
    return fail(hipcompErrorInvalidValue, fn,
                "Batch too large for the built-in scratch: " + std::to_string(batch_size) + " chunks of up to "
                    + std::to_string(max_uncompressed_chunk_bytes) + " bytes need " + std::to_string(required)
                    + " bytes of hash tables, the library holds " + std::to_string(B200COMP_LZ4_NOALLOC_SCRATCH_BYTES)
                    + " (at most " + std::to_string(B200COMP_LZ4_NOALLOC_MAX_BATCH)
                    + " chunks above 64 KiB per call; use hipcompBatchedLZ4CompressAsync with a workspace).");
  void* scratch = nullptr;
  const cudaError_t e = noalloc_scratch(&scratch);
  if (e != cudaSuccess)
    return fail(hipcompErrorCudaError, fn, std::string("cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
  return hipcompBatchedLZ4CompressAsync(device_uncompressed_ptrs, device_uncompressed_bytes,
                                        max_uncompressed_chunk_bytes, batch_size, scratch,
                                        (size_t)B200COMP_LZ4_NOALLOC_SCRATCH_BYTES, device_compressed_ptrs,
                                        device_compressed_bytes, format_opts, stream);
}
