// b200comp — batched Snappy raw-block codec for sm_100a.
//
// C ABI replaces reference src/lowlevel/SnappyBatch.cpp:84-245.
//
// Encoders.  Chunks up to 64 KiB (every headline configuration) go to the team encoder of
// lz_team_encode.cuh with the Snappy Emit policy below: one CTA per chunk, chunk and hash table
// resident in shared memory.  Its streams are plain Snappy streams with a ratio above the
// reference's, but not the reference's bytes.  Larger chunks (and callers that declare no chunk
// size) keep the warp encoder below, which reproduces the reference's match policy
// (src/snappy/compression.hiph:190-385: 12-bit hash, 32-position windows, nearest-lower-lane twin,
// uint16 table, <= 64-byte copies) byte for byte - one warp does search and emission back to back
// where the reference ping-pongs two warps through __syncthreads().
// The decoder is one warp per chunk over the element grammar of
// src/snappy/decompression_decode.hiph:72-150 (the reference uses a 3-warp prefetch / decode /
// process pipeline with nanosleep spin queues); see lz4.cu and lz_common.cuh for the scheme.
#include "lz_common.cuh"
#include "lz_team_encode.cuh"

#pragma GCC visibility push(default)
#include "hipcomp/snappy.h"
#pragma GCC visibility pop

namespace b200comp {
namespace snappy {

constexpr uint32_t kHashBits = 12;
constexpr uint32_t kMaxLiteral = 256;
constexpr uint32_t kMaxCopyDistance = 32768;
constexpr int kCompWarps = 4;
constexpr int kDecWarps = 4;

__device__ __forceinline__ uint32_t snap_hash(uint32_t v) { return (v * 0x102A6Bu) >> (32 - kHashBits); }

// literal element header + bytes (compression.hiph:73-117)
__device__ __forceinline__ uint32_t put_literal(uint8_t* dst, uint32_t o, const uint8_t* src, uint32_t len)
{
  const uint32_t l1 = len - 1;
  uint32_t hdr;
  if (l1 < 60) hdr = 1;
  else if (l1 <= 0xff) hdr = 2;
  else if (l1 <= 0xffff) hdr = 3;
  else if (l1 <= 0xffffff) hdr = 4;
  else hdr = 5;
  if (lane_id() == 0) {
    if (hdr == 1) {
      dst[o] = (uint8_t)(l1 << 2);
    } else {
      dst[o] = (uint8_t)((58 + hdr) << 2);
      for (uint32_t i = 0; i + 1 < hdr; ++i) dst[o + 1 + i] = (uint8_t)(l1 >> (8 * i));
    }
  }
  o += hdr;
  warp_copy_ro(dst + o, src, len);
  return o + len;
}

// copy element (compression.hiph:129-151)
__device__ __forceinline__ uint32_t put_copy(uint8_t* dst, uint32_t o, uint32_t len, uint32_t dist)
{
  if (len < 12 && dist < 2048) {
    if (lane_id() == 0) {
      dst[o] = (uint8_t)(((dist & 0x700) >> 3) | ((len - 4) << 2) | 0x01);
      dst[o + 1] = (uint8_t)dist;
    }
    return o + 2;
  }
  if (lane_id() == 0) {
    dst[o] = (uint8_t)(((len - 1) << 2) | 0x2);
    dst[o + 1] = (uint8_t)dist;
    dst[o + 2] = (uint8_t)(dist >> 8);
  }
  return o + 3;
}

// One warp per chunk, an 8 KiB hash map per warp in (dynamic) shared memory.  WARPS = warps per CTA:
// 4 (six CTAs per SM) or 28 (one CTA per SM, 224 KiB of maps): the launcher takes the shape that
// needs fewer waves over the batch, e.g. 4096 chunks (C1) fit 148 x 28 = 4144 warps in one wave.
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32, WARPS == 4 ? 6 : 1)
compress_kernel(const void* const* in_ptrs, const size_t* in_bytes, void* const* out_ptrs, size_t* out_bytes,
                size_t batch)
{
  extern __shared__ __align__(16) uint16_t maps[];
  const uint32_t lane = (uint32_t)lane_id();
  const int warp = threadIdx.x >> 5;
  uint16_t* map = maps + (size_t)warp * (1 << kHashBits);
  const unsigned lt = lanemask_lt();
  const size_t stride = (size_t)gridDim.x * WARPS;

  for (size_t chunk = (size_t)blockIdx.x * WARPS + warp; chunk < batch; chunk += stride) {
    const uint8_t* src = static_cast<const uint8_t*>(in_ptrs[chunk]);
    uint8_t* dst = static_cast<uint8_t*>(out_ptrs[chunk]);
    const uint32_t len = (uint32_t)in_bytes[chunk];
    {
      uint32_t* m32 = reinterpret_cast<uint32_t*>(map);
      for (uint32_t i = lane; i < (1u << kHashBits) / 2; i += 32) m32[i] = 0;
    }
    uint32_t o = 0;
    if (lane == 0) {
      uint32_t v = len;
      while (v > 0x7f) { dst[o++] = (uint8_t)(v | 0x80); v >>= 7; }
      dst[o++] = (uint8_t)v;
    }
    o = __shfl_sync(kFull, o, 0);
    __syncwarp();

    uint32_t pos = 0;
    while (pos < len) {
      // ---- search round: up to 8 windows of 32 positions (FindFourByteMatch)
      uint32_t p = pos;
      const uint32_t maxpos = pos + kMaxLiteral - 31;
      uint32_t literal_cnt;
      uint32_t copy_len = 0, dist = 0;
      do {
        const bool valid4 = p + lane + 4 <= len;
        const uint32_t data = valid4 ? ld_unaligned32_ro(src + p + lane) : 0u;
        const uint32_t hash = valid4 ? snap_hash(data) : 0u;
        const unsigned group = __match_any_sync(kFull, hash);
        const unsigned lower = group & lt;
        const uint32_t lower_lane = lower ? 31u - (uint32_t)__clz((int)lower) : lane;
        const uint32_t lower_data = __shfl_sync(kFull, data, (int)lower_lane);
        uint32_t offset = p + lane;
        bool match = false;
        if (valid4) {
          if (lower && lower_data == data) {
            match = true;
            offset = p + lower_lane;
          } else {
            offset = (p & ~0xffffu) | map[hash];
            if (offset >= p) offset = offset >= 0x10000u ? offset - 0x10000u : p;
            match = offset < p && offset + kMaxCopyDistance >= p + lane && ld_unaligned32_ro(src + offset) == data;
          }
        }
        const unsigned mm = __ballot_sync(kFull, match);
        if (mm) {
          literal_cnt = (uint32_t)__ffs((int)mm) - 1u;
          dist = p + literal_cnt - __shfl_sync(kFull, offset, (int)literal_cnt);
          copy_len = 4;
        } else {
          literal_cnt = 32;
        }
        // table update: highest lane of each equal-hash group among lanes <= literal_cnt
        const unsigned keep = literal_cnt >= 31 ? kFull : ((2u << literal_cnt) - 1u);
        const unsigned g = valid4 ? (group & keep) : 0u;
        if (lane <= literal_cnt && g && 31u - (uint32_t)__clz((int)g) == lane) map[hash] = (uint16_t)(p + lane);
        __syncwarp();
        p += literal_cnt;
      } while (literal_cnt == 32 && p < maxpos);
      const uint32_t lit = min(p, len) - pos;
      // ---- extend the 4-byte match by up to 60 bytes (Match60)
      if (copy_len) {
        const uint32_t mpos = pos + lit + 4;
        const uint32_t lim = min(len - mpos, 60u);
        const bool ne0 = lane >= lim || __ldg(src + mpos + lane) != __ldg(src + mpos - dist + lane);
        const unsigned m0 = __ballot_sync(kFull, ne0);
        if (m0) {
          copy_len += (uint32_t)__ffs((int)m0) - 1u;
        } else {
          const bool ne1 = 32 + lane >= lim || __ldg(src + mpos + 32 + lane) != __ldg(src + mpos - dist + 32 + lane);
          const unsigned m1 = __ballot_sync(kFull, ne1);
          copy_len += 31u + (uint32_t)__ffs((int)m1);
        }
      }
      // ---- emit
      if (lit) o = put_literal(dst, o, src + pos, lit);
      pos += lit;
      if (copy_len) o = put_copy(dst, o, copy_len, dist);
      pos += copy_len;
    }
    if (lane == 0) out_bytes[chunk] = o;
    __syncwarp();
  }
}

// ------------------------------------------------------------------ decode

// varint preamble; bytes past the end read as 0, fifth byte >= 8 is an error
// (SnappyBatchKernels.hip:84-134, decompression.hiph:71-104)
__device__ __forceinline__ bool read_preamble(const uint8_t* comp, uint32_t comp_len, uint32_t& cur, uint32_t& size)
{
  size = 0;
  cur = 0;
#pragma unroll
  for (int shift = 0; shift <= 28; shift += 7) {
    const uint32_t c = cur < comp_len ? (uint32_t)__ldg(comp + cur) : 0u;
    if (cur < comp_len) ++cur;
    if (shift == 28) {
      if (c >= 8) return false;
      size |= c << 28;
      break;
    }
    size |= (c & 0x7f) << shift;
    if (!(c & 0x80)) break;
  }
  return true;
}

// The decoder (same scheme as lz4::decompress_window_kernel): one warp per chunk.  A round stages
// 512 compressed bytes in shared memory and computes, for every byte position, how far the next
// element starts if an element started there (tag kinds folded with SIMD-in-word arithmetic), so
// the serial walk is one shared-memory byte load per element.  Lane k then decodes element k from
// the window; validation (offset within the output so far, length within the bytes left) runs in
// parallel over the round after a prefix sum, and an invalid element cuts the round in front of
// it: the elements before it are still produced, as the sequential decoder does.  Literals with
// an extended length (>= 61 bytes) and elements that do not fit a fresh window run one at a time
// from global memory with vector copies.
constexpr int kWin = 512;
constexpr uint32_t kWinKeep = 160; // a staged window is reused while at least this much of it lies ahead

__global__ void __launch_bounds__(kDecWarps * 32)
decompress_window_kernel(const void* const* comp_ptrs, const size_t* comp_bytes, const size_t* out_caps,
                         void* const* out_ptrs, size_t* actual, hipcompStatus_t* statuses, size_t batch)
{
  __shared__ __align__(16) uint8_t s_win[kDecWarps][kWin + 32];
  __shared__ __align__(16) uint8_t s_step[kDecWarps][kWin + 64]; // + 64 zero bytes: a walk that leaves the window stops
  __shared__ uint32_t s_own[kDecWarps][kOwnWords];           // start bitmap, bases and stage of the round executor
  __shared__ uint16_t s_idx[kDecWarps][32];                  // element positions found by the walk
  const uint32_t lane = (uint32_t)lane_id();
  const int warp = threadIdx.x >> 5;
  uint8_t* win = s_win[warp];
  uint8_t* step = s_step[warp];
  reinterpret_cast<uint16_t*>(step + kWin)[lane_id()] = 0; // the pad behind the table (never written again)
  const size_t stride = (size_t)gridDim.x * kDecWarps;
  for (size_t chunk = (size_t)blockIdx.x * kDecWarps + warp; chunk < batch; chunk += stride) {
    const uint8_t* comp = static_cast<const uint8_t*>(comp_ptrs[chunk]);
    uint8_t* out = static_cast<uint8_t*>(out_ptrs[chunk]);
    const uint32_t comp_len = (uint32_t)comp_bytes[chunk];
    const size_t cap64 = out_caps[chunk];
    const uintptr_t cbase = reinterpret_cast<uintptr_t>(comp);
    uint32_t size = 0, cur = 0, left = 0;
    bool ok = comp_len != 0 && comp != nullptr;
    if (ok) ok = read_preamble(comp, comp_len, cur, size);
    if (ok) {
      const size_t cap = cap64 == 0 ? size : cap64;
      if ((cur >= comp_len && size != 0) || size > cap) ok = false;
    }
    if (!ok) size = 0;
    left = size;
    uint32_t dst = 0;
    // the staged window [wpos, wpos + wvalid) and its step table stay valid across rounds: a round consumes
    // ~120 of the 512 bytes, so three rounds out of four start without a load from global memory
    uint32_t wpos = 0, wvalid = 0;
    bool staged = false;
    while (ok && left > 0) {
      if (cur >= comp_len) { ok = false; break; }
      // reuse while a good part of a round still lies inside the window (or the window ends with the stream)
      uint32_t shift = cur - wpos; // where the walk starts inside the window (wpos may lie up to 15 bytes below 0: wrap-safe)
      const bool reuse = staged && shift < wvalid && (shift + kWinKeep <= wvalid || wpos + wvalid == comp_len);
      const bool fresh = !reuse;
      if (fresh) {
      // ---- stage the window
      shift = (uint32_t)((cbase + cur) & 15);
      wpos = cur - shift;
      wvalid = min((uint32_t)kWin, comp_len - wpos);
      staged = true;
      __syncwarp(); // the previous round's readers of the window are done
      uint4 v = make_uint4(0, 0, 0, 0);
      if (16u * lane < wvalid) v = ld_block16_guarded(cbase, comp_len, cbase + cur - shift + 16u * lane, win + 16 * lane);
      *reinterpret_cast<uint4*>(win + 16 * lane) = v;
      {
        // step[p] for a tag at p: literal of 1..60 bytes -> 1 + length; copy kinds 1/2/3 -> 2/3/5;
        // 0 for a literal with an extended length
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
        uint32_t r[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t k1 = w[j] & 0x01010101u, k2 = (w[j] >> 1) & 0x01010101u;
          const uint32_t len6 = (w[j] >> 2) & 0x3f3f3f3fu;
          const uint32_t is_copy = (k1 | k2) * 0xffu;
          const uint32_t ext = ((len6 + 0x04040404u) >> 6) & 0x01010101u; // len6 >= 60
          const uint32_t lit_step = (len6 + 0x02020202u) & ~(ext * 0xffu);
          const uint32_t copy_step = 0x01010101u + k1 + 2u * k2 + (k1 & k2);
          r[j] = (copy_step & is_copy) | (lit_step & ~is_copy);
        }
        *reinterpret_cast<uint4*>(step + 16 * lane) = make_uint4(r[0], r[1], r[2], r[3]);
        // an element that does not end inside the window (or starts behind it) gets step 0 as well, so that the
        // walk below tests one thing; only the lanes within 80 bytes of the window's end have such entries
        if (16u * lane + 80u > wvalid) {
#pragma unroll 1
          for (uint32_t b = 0; b < 16u; ++b) {
            const uint32_t p = 16u * lane + b;
            if (p + step[p] > wvalid) step[p] = 0;
          }
        }
      }
      } // fresh
      __syncwarp();

      // ---- walk the element chain: one table byte per element, 0 = stop (extended literal, window end)
      uint32_t idx = shift, k = 0;
      uint16_t* idxs = s_idx[warp];
#pragma unroll 8
      for (; k < 32u; ++k) {
        const uint32_t s = step[idx]; // idx < kWin + 61: inside the table or its zero pad
        if (s == 0u) break;
        idxs[k] = (uint16_t)idx;
        idx += s;
      }
      __syncwarp();
      const uint32_t my_idx = lane < k ? (uint32_t)idxs[lane] : 0u;
      if (k == 0u && !fresh) { // the element did not fit what is left of a reused window: stage at it
        staged = false;
        continue;
      }
      if (k == 0u) {
        // one element straight from global memory: extended literal, or an element cut by the window end
        uint32_t blen, offset = 0;
        const uint32_t b0 = __ldg(comp + cur);
        if (b0 & 3) {
          if ((b0 & 3) == 1) {
            if (cur + 2 > comp_len) { ok = false; break; }
            offset = ((b0 & 0xe0) << 3) | __ldg(comp + cur + 1);
            blen = ((b0 >> 2) & 7) + 4;
            cur += 2;
          } else if ((b0 & 3) == 2) {
            if (cur + 3 > comp_len) { ok = false; break; }
            offset = (uint32_t)__ldg(comp + cur + 1) | ((uint32_t)__ldg(comp + cur + 2) << 8);
            blen = (b0 >> 2) + 1;
            cur += 3;
          } else {
            if (cur + 5 > comp_len) { ok = false; break; }
            offset = (uint32_t)__ldg(comp + cur + 1) | ((uint32_t)__ldg(comp + cur + 2) << 8)
                     | ((uint32_t)__ldg(comp + cur + 3) << 16) | ((uint32_t)__ldg(comp + cur + 4) << 24);
            blen = (b0 >> 2) + 1;
            cur += 5;
          }
          if (offset == 0 || offset > dst || left < blen) { ok = false; break; }
          __syncwarp();
          group_copy_match<32>(out + dst, offset, blen, lane, kFull);
        } else {
          blen = b0 >> 2;
          cur += 1;
          if (blen >= 60) {
            const uint32_t nb = blen - 59;
            if (cur + nb > comp_len) { ok = false; break; }
            blen = 0;
            for (uint32_t i = 0; i < nb; ++i) blen |= (uint32_t)__ldg(comp + cur + i) << (8 * i);
            cur += nb;
          }
          if (blen == 0xffffffffu) { ok = false; break; }
          blen += 1;
          if (left < blen || blen > comp_len - cur) { ok = false; break; }
          group_copy_ro<32>(out + dst, comp + cur, blen, lane);
          cur += blen;
        }
        dst += blen;
        left -= blen;
        __syncwarp();
        continue;
      }
      // ---- lane j decodes element j
      uint32_t my_len = 0, my_off = 0, my_lp = 0; // my_off == 0: literal
      bool zero_off = false;
      if (lane < k) {
        const uint32_t tag = win[my_idx], kind = tag & 3u;
        if (kind == 0u) {
          my_len = (tag >> 2) + 1u;
          my_lp = wpos + my_idx + 1u;
        } else {
          if (kind == 1u) {
            my_off = ((tag & 0xe0u) << 3) | win[my_idx + 1u];
            my_len = ((tag >> 2) & 7u) + 4u;
          } else {
            my_off = (uint32_t)win[my_idx + 1u] | ((uint32_t)win[my_idx + 2u] << 8);
            if (kind == 3u) my_off |= ((uint32_t)win[my_idx + 3u] << 16) | ((uint32_t)win[my_idx + 4u] << 24);
            my_len = (tag >> 2) + 1u;
          }
          zero_off = my_off == 0u;
        }
      }
      // ---- validate in stream order: an element must fit the bytes left and a copy must start
      //      inside the output produced so far; elements past the end of the output are ignored
      uint32_t total;
      const uint32_t rel = warp_excl_sum_sat(my_len, total);
      const bool past_end = lane < k && rel >= left;            // output already complete before this element
      const bool bad = lane < k && !past_end && (my_len > left - rel || zero_off || my_off > dst + rel);
      const unsigned stop = __ballot_sync(kFull, bad || past_end || lane >= k);
      const uint32_t nrun = stop ? (uint32_t)__ffs((int)stop) - 1u : 32u; // elements that run
      const unsigned bad_mask = __ballot_sync(kFull, bad);
      const bool failed = nrun < 32u && ((bad_mask >> nrun) & 1u);
      if (lane >= nrun) my_len = 0;
      const uint32_t run_total = __shfl_sync(kFull, rel, (int)(nrun & 31u)) ;
      const uint32_t produced = nrun == 32u ? total : run_total;
      const uint32_t pos = dst + rel;
      const uint32_t seq_end = pos + my_len;
      const uint32_t lit_end = my_off == 0 ? seq_end : pos;
      if (!execute_round_owned(out, win, s_own[warp], dst, dst + produced, pos, lit_end, seq_end, my_lp - wpos - pos, my_off))
        execute_round(out, comp, dst, dst + produced, pos, lit_end, seq_end, my_lp - pos, my_off);
      dst += produced;
      left -= produced;
      if (failed) { ok = false; break; }
      // stream position after the last element that ran
      const uint32_t nidx = __shfl_sync(kFull, my_idx, (int)(nrun & 31u));
      cur = wpos + (nrun == k ? idx : nidx);
      __syncwarp();
    }
    if (lane == 0) {
      if (actual) actual[chunk] = size - left;
      if (statuses) statuses[chunk] = ok ? hipcompSuccess : hipcompErrorCannotDecompress;
    }
    __syncwarp();
  }
}

__global__ void size_kernel(const void* const* comp_ptrs, const size_t* comp_bytes, size_t* sizes, size_t batch)
{
  for (size_t chunk = (size_t)blockIdx.x * blockDim.x + threadIdx.x; chunk < batch;
       chunk += (size_t)gridDim.x * blockDim.x) {
    const uint8_t* comp = static_cast<const uint8_t*>(comp_ptrs[chunk]);
    const uint32_t comp_len = (uint32_t)comp_bytes[chunk];
    uint32_t cur, size = 0;
    if (comp_len == 0 || comp == nullptr || !read_preamble(comp, comp_len, cur, size)) size = 0;
    sizes[chunk] = size;
  }
}

// ---- team encoder (lz_team_encode.cuh) for chunks <= 64 KiB: how a Snappy sequence is sized and written
//      (elements: src/snappy/symbol.hiph:43-70).  oracle/lz_team_model.c states the same sequentially.
constexpr size_t kTeamMaxChunk = 65536;

struct TeamEmit {
  static constexpr bool kNeedDist = true;
  static __device__ __forceinline__ uint32_t empty_stream(uint8_t* gout)
  {
    gout[0] = 0; // varint(0)
    return 1;
  }
  // preamble: the uncompressed length as a varint
  static __device__ __forceinline__ uint32_t preamble(uint8_t* gout, uint32_t n, uint32_t lane)
  {
    const uint32_t nb = n < 0x80u ? 1u : (n < 0x4000u ? 2u : 3u);
    if (lane < nb) gout[lane] = (uint8_t)(((n >> (7u * lane)) & 0x7fu) | (lane + 1u < nb ? 0x80u : 0u));
    return nb;
  }
  static __device__ __forceinline__ uint32_t lit_bytes(uint32_t nlit)
  {
    return nlit ? nlit + 1u + (nlit > 60u ? 1u : 0u) + (nlit > 256u ? 1u : 0u) : 0u;
  }
  // copies of 64 while 68 or more remain, one of 60 when more than 64 remain, then the rest (>= 4)
  static __device__ __forceinline__ uint32_t match_bytes(uint32_t len, uint32_t d)
  {
    uint32_t bytes = 0;
    if (len >= 68u) {
      const uint32_t k = (len - 68u) / 64u + 1u;
      bytes = 3u * k;
      len -= 64u * k;
    }
    if (len > 64u) {
      bytes += 3u;
      len -= 60u;
    }
    return bytes + ((len < 12u && d < 2048u) ? 2u : 3u);
  }
  static __device__ __forceinline__ uint8_t* put_lit_header(uint8_t* o, uint32_t nlit, uint32_t)
  {
    if (nlit == 0u) return o;
    const uint32_t l1 = nlit - 1u;
    if (l1 < 60u) {
      *o++ = (uint8_t)(l1 << 2);
    } else if (l1 <= 0xffu) {
      *o++ = (uint8_t)(60u << 2);
      *o++ = (uint8_t)l1;
    } else {
      *o++ = (uint8_t)(61u << 2);
      *o++ = (uint8_t)l1;
      *o++ = (uint8_t)(l1 >> 8);
    }
    return o;
  }
  static __device__ __forceinline__ uint8_t* put_match(uint8_t* o, uint32_t d, uint32_t len)
  {
    while (len) {
      uint32_t piece = len;
      if (len >= 68u) piece = 64u;
      else if (len > 64u) piece = 60u;
      if (piece < 12u && d < 2048u) {
        *o++ = (uint8_t)(((d & 0x700u) >> 3) | ((piece - 4u) << 2) | 0x01u);
        *o++ = (uint8_t)d;
      } else {
        *o++ = (uint8_t)(((piece - 1u) << 2) | 0x2u);
        *o++ = (uint8_t)d;
        *o++ = (uint8_t)(d >> 8);
      }
      len -= piece;
    }
    return o;
  }
  // the copies of a long match (len > team::kLongMatch), written by the whole warp: 64-byte copies, one per lane
  // and round, then the rest by one lane
  static __device__ __forceinline__ void put_match_warp(uint8_t* o, uint32_t d, uint32_t len, uint32_t lane)
  {
    const uint32_t k = (len - 68u) / 64u + 1u;
    for (uint32_t i = lane; i < k; i += 32u) {
      o[3u * i] = (uint8_t)(((64u - 1u) << 2) | 0x2u);
      o[3u * i + 1u] = (uint8_t)d;
      o[3u * i + 2u] = (uint8_t)(d >> 8);
    }
    if (lane == 0) put_match(o + 3u * k, d, len - 64u * k);
  }
  static __device__ __forceinline__ uint32_t put_tail_header(uint8_t* o, uint32_t nlit, uint32_t lane)
  {
    if (lane == 0) put_lit_header(o, nlit, 0);
    return lit_bytes(nlit) - nlit;
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

} // namespace snappy
} // namespace b200comp

using namespace b200comp;

extern "C" {

hipcompStatus_t hipcompBatchedSnappyDecompressGetTempSize(size_t, size_t, size_t* temp_bytes)
{
  B200_CHECK_NOT_NULL(temp_bytes, "hipcompBatchedSnappyDecompressGetTempSize()");
  *temp_bytes = 0;
  return hipcompSuccess;
}

hipcompStatus_t hipcompBatchedSnappyGetDecompressSizeAsync(const void* const* device_compressed_ptrs,
                                                           const size_t* device_compressed_bytes,
                                                           size_t* device_uncompressed_bytes, size_t batch_size,
                                                           hipStream_t stream)
{
  static const char* fn = "hipcompBatchedSnappyGetDecompressSizeAsync()";
  B200_TRACE(fn);
  B200_CHECK_NOT_NULL(device_compressed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  if (batch_size == 0) return hipcompSuccess;
  const size_t need = (batch_size + 127) / 128;
  const size_t cap = (size_t)sm_count() * 16;
  snappy::size_kernel<<<(int)(need < cap ? need : cap), 128, 0, stream>>>(
      device_compressed_ptrs, device_compressed_bytes, device_uncompressed_bytes, batch_size);
  return check_launch(fn);
}

hipcompStatus_t hipcompBatchedSnappyDecompressAsync(const void* const* device_compresed_ptrs,
                                                    const size_t* device_compressed_bytes,
                                                    const size_t* device_uncompressed_bytes,
                                                    size_t* device_actual_uncompressed_bytes, size_t batch_size,
                                                    void* const, const size_t, void* const* device_uncompressed_ptr,
                                                    hipcompStatus_t* device_statuses, hipStream_t stream)
{
  static const char* fn = "hipcompBatchedSnappyDecompressAsync()";
  B200_TRACE(fn);
  B200_CHECK_NOT_NULL(device_compresed_ptrs, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_ptr, fn);
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_DEVICE_PTR(device_compresed_ptrs, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_ptr, fn);
  const size_t need = (batch_size + snappy::kDecWarps - 1) / snappy::kDecWarps;
  const size_t cap = (size_t)sm_count() * 16;
  snappy::decompress_window_kernel<<<(int)(need < cap ? need : cap), snappy::kDecWarps * 32, 0, stream>>>(
      device_compresed_ptrs, device_compressed_bytes, device_uncompressed_bytes, device_uncompressed_ptr,
      device_actual_uncompressed_bytes, device_statuses, batch_size);
  return check_launch(fn);
}

hipcompStatus_t hipcompBatchedSnappyCompressGetTempSize(size_t, size_t, hipcompBatchedSnappyOpts_t, size_t* temp_bytes)
{
  B200_CHECK_NOT_NULL(temp_bytes, "hipcompBatchedSnappyCompressGetTempSize()");
  *temp_bytes = 0;
  return hipcompSuccess;
}

hipcompStatus_t hipcompBatchedSnappyCompressGetMaxOutputChunkSize(size_t max_chunk_size, hipcompBatchedSnappyOpts_t,
                                                                  size_t* max_compressed_size)
{
  B200_CHECK_NOT_NULL(max_compressed_size, "hipcompBatchedSnappyCompressGetOutputSize()");
  *max_compressed_size = 32 + max_chunk_size + max_chunk_size / 6;
  return hipcompSuccess;
}

hipcompStatus_t hipcompBatchedSnappyCompressAsync(const void* const* device_uncompressed_ptr,
                                                  const size_t* device_uncompressed_bytes,
                                                  size_t max_uncompressed_chunk_bytes, size_t batch_size,
                                                  void*, size_t, void* const* device_compressed_ptr,
                                                  size_t* device_compressed_bytes, hipcompBatchedSnappyOpts_t,
                                                  hipStream_t stream)
{
  static const char* fn = "hipcompBatchedSnappyCompressAsync()";
  B200_TRACE(fn);
  B200_CHECK_NOT_NULL(device_uncompressed_ptr, fn);
  B200_CHECK_NOT_NULL(device_uncompressed_bytes, fn);
  B200_CHECK_NOT_NULL(device_compressed_ptr, fn);
  B200_CHECK_NOT_NULL(device_compressed_bytes, fn);
  if (batch_size == 0) return hipcompSuccess;
  B200_CHECK_DEVICE_PTR(device_uncompressed_ptr, fn);
  B200_CHECK_DEVICE_PTR(device_uncompressed_bytes, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_ptr, fn);
  B200_CHECK_DEVICE_PTR(device_compressed_bytes, fn);
  if (max_uncompressed_chunk_bytes != 0 && max_uncompressed_chunk_bytes <= snappy::kTeamMaxChunk) {
    snappy::launch_team(device_uncompressed_ptr, device_uncompressed_bytes, max_uncompressed_chunk_bytes, batch_size,
                        device_compressed_ptr, device_compressed_bytes, stream);
    return check_launch(fn);
  }
  // larger (or undeclared) chunk sizes: the reference-policy warp encoder
  const size_t sms = (size_t)sm_count();
  constexpr size_t kMapBytes = sizeof(uint16_t) << snappy::kHashBits;
  // waves over the batch: 24 warps per SM as 6 CTAs of 4, or 28 as one CTA
  const size_t waves4 = (batch_size + sms * 24 - 1) / (sms * 24), waves28 = (batch_size + sms * 28 - 1) / (sms * 28);
  const bool big = waves28 < waves4;
  if (big) {
    constexpr int W = 28;
    allow_smem(snappy::compress_kernel<W>, (int)(W * kMapBytes));
    const size_t need = (batch_size + W - 1) / W;
    snappy::compress_kernel<W><<<(int)(need < sms ? need : sms), W * 32, W * kMapBytes, stream>>>(
        device_uncompressed_ptr, device_uncompressed_bytes, device_compressed_ptr, device_compressed_bytes, batch_size);
  } else {
    constexpr int W = snappy::kCompWarps;
    const size_t need = (batch_size + W - 1) / W;
    const size_t cap = sms * 6;
    snappy::compress_kernel<W><<<(int)(need < cap ? need : cap), W * 32, W * kMapBytes, stream>>>(
        device_uncompressed_ptr, device_uncompressed_bytes, device_compressed_ptr, device_compressed_bytes, batch_size);
  }
  return check_launch(fn);
}

} // extern "C"
