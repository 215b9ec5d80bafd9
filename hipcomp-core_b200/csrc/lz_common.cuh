// b200comp — what the LZ4 and Snappy kernels share: warp-level byte movers, the generic round
// executor (execute_round) and the decoders' fast round executor (execute_round_owned).
#pragma once

#include "common.cuh"

namespace b200comp {

// 4 bytes at an arbitrary byte address of read-only data: two aligned word
// loads + funnel shift.  Touches at most the aligned words that contain
// [p, p+4), so it never reads a byte outside the 4-byte-aligned hull of the
// requested range.
__device__ __forceinline__ uint32_t ld_unaligned32_ro(const uint8_t* p)
{
  const uintptr_t a = reinterpret_cast<uintptr_t>(p);
  const uint32_t* w = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
  const uint32_t sh = (uint32_t)(a & 3) * 8;
  const uint32_t lo = __ldg(w);
  if (sh == 0) return lo;
  const uint32_t hi = __ldg(w + 1);
  return __funnelshift_r(lo, hi, sh);
}

// The 16-byte aligned block at `block` of a stream [base, base + len), for the window decoders (they stage whole
// aligned blocks; the first and the last one of a stream may stick out of the caller's buffer): one vector load
// when the block lies inside the stream, else byte loads of the bytes that do, through the lane's own 16 bytes
// of the shared-memory window `slot` (the rest reads as 0).  The edge path is a rolled loop: two lanes per chunk
// ever take it, and it must not cost the common path registers.
__device__ __forceinline__ uint4 ld_block16_guarded(uintptr_t base, uint32_t len, uintptr_t block, uint8_t* slot)
{
  if (block >= base && block + 16 <= base + len) return __ldg(reinterpret_cast<const uint4*>(block));
  *reinterpret_cast<uint4*>(slot) = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll 1
  for (uint32_t i = 0; i < 16u; ++i) {
    const uintptr_t a = block + i;
    if (a >= base && a < base + len) slot[i] = __ldg(reinterpret_cast<const uint8_t*>(a));
  }
  return *reinterpret_cast<const uint4*>(slot);
}

// Warp-cooperative copy of n bytes from read-only memory to `dst`.
// Short runs: one byte per lane.  Long runs: byte prologue up to a 16-byte
// aligned destination, then 16-byte stores fed by aligned word loads
// (funnel-shifted when source and destination are mutually misaligned).
__device__ __forceinline__ void warp_copy_ro(uint8_t* dst, const uint8_t* src, uint32_t n)
{
  const uint32_t lane = (uint32_t)lane_id();
  if (n < 96) {
    for (uint32_t i = lane; i < n; i += 32) dst[i] = __ldg(src + i);
    return;
  }
  const uint32_t head = (uint32_t)((16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15);
  if (lane < head) dst[lane] = __ldg(src + lane);
  dst += head;
  src += head;
  n -= head;
  const uint32_t vecs = n >> 4;
  const uintptr_t sa = reinterpret_cast<uintptr_t>(src);
  const uint32_t* sw = reinterpret_cast<const uint32_t*>(sa & ~(uintptr_t)3);
  const uint32_t sh = (uint32_t)(sa & 3) * 8;
  uint4* dv = reinterpret_cast<uint4*>(dst);
  if (sh == 0) {
    for (uint32_t v = lane; v < vecs; v += 32) {
      const uint32_t* s = sw + 4 * v;
      uint4 r;
      r.x = __ldg(s); r.y = __ldg(s + 1); r.z = __ldg(s + 2); r.w = __ldg(s + 3);
      dv[v] = r;
    }
  } else {
    for (uint32_t v = lane; v < vecs; v += 32) {
      const uint32_t* s = sw + 4 * v;
      const uint32_t a = __ldg(s), b = __ldg(s + 1), c = __ldg(s + 2), d = __ldg(s + 3), e = __ldg(s + 4);
      uint4 r;
      r.x = __funnelshift_r(a, b, sh);
      r.y = __funnelshift_r(b, c, sh);
      r.z = __funnelshift_r(c, d, sh);
      r.w = __funnelshift_r(d, e, sh);
      dv[v] = r;
    }
  }
  for (uint32_t i = (vecs << 4) + lane; i < n; i += 32) dst[i] = __ldg(src + i);
}

// Warp-cooperative LZ77 match copy inside the output buffer:
// dst[i] = dst[i - off] for i in [0, len), semantics of a sequential byte copy.
// Callers have made all earlier writes to the buffer visible (__syncwarp()).
__device__ __forceinline__ void warp_copy_match(uint8_t* dst, uint32_t off, uint32_t len)
{
  const uint32_t lane = (uint32_t)lane_id();
  const uint8_t* src = dst - off;
  if (off >= len) {
    // no overlap: plain copy (data was written by this warp: coherent loads)
    for (uint32_t i = lane; i < len; i += 32) dst[i] = src[i];
  } else if (off >= 32) {
    // overlap further than one warp-wide step: steps are ordered by a barrier
    for (uint32_t base = 0; base < len; base += 32) {
      const uint32_t i = base + lane;
      uint8_t v = 0;
      if (i < len) v = src[i];
      if (i < len) dst[i] = v;
      __syncwarp();
    }
  } else {
    // periodic pattern of `off` bytes that all precede dst
    for (uint32_t i = lane; i < len; i += 32) dst[i] = src[i % off];
  }
}

} // namespace b200comp

// ---------------------------------------------------------------- sub-warp groups
//
// The decoders give each chunk to a group of G lanes (G = 4..32, a power of two), so
// that one warp instruction advances 32/G chunks at once: typical LZ sequences move
// 5-20 bytes, which leaves most of a 32-lane copy idle.  Everything below is
// synchronised with the group's own lane mask only.

namespace b200comp {

template <int G>
__device__ __forceinline__ unsigned group_mask()
{
  if (G == 32) return kFull;
  return ((1u << (G & 31)) - 1u) << ((unsigned)lane_id() & ~(unsigned)(G - 1));
}

// n bytes of read-only memory -> dst, by the G lanes of a group (sub = lane % G).
template <int G>
__device__ __forceinline__ void group_copy_ro(uint8_t* dst, const uint8_t* src, uint32_t n, uint32_t sub)
{
  if (n < 16u * G) {
    for (uint32_t i = sub; i < n; i += G) dst[i] = __ldg(src + i);
    return;
  }
  const uint32_t head = (uint32_t)((16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15);
  for (uint32_t i = sub; i < head; i += G) dst[i] = __ldg(src + i);
  dst += head;
  src += head;
  n -= head;
  const uint32_t vecs = n >> 4;
  const uintptr_t sa = reinterpret_cast<uintptr_t>(src);
  const uint32_t* sw = reinterpret_cast<const uint32_t*>(sa & ~(uintptr_t)3);
  const uint32_t sh = (uint32_t)(sa & 3) * 8;
  uint4* dv = reinterpret_cast<uint4*>(dst);
  if (sh == 0) {
    for (uint32_t v = sub; v < vecs; v += G) {
      const uint32_t* s = sw + 4 * v;
      uint4 r;
      r.x = __ldg(s); r.y = __ldg(s + 1); r.z = __ldg(s + 2); r.w = __ldg(s + 3);
      dv[v] = r;
    }
  } else {
    for (uint32_t v = sub; v < vecs; v += G) {
      const uint32_t* s = sw + 4 * v;
      const uint32_t a = __ldg(s), b = __ldg(s + 1), c = __ldg(s + 2), d = __ldg(s + 3), e = __ldg(s + 4);
      uint4 r;
      r.x = __funnelshift_r(a, b, sh);
      r.y = __funnelshift_r(b, c, sh);
      r.z = __funnelshift_r(c, d, sh);
      r.w = __funnelshift_r(d, e, sh);
      dv[v] = r;
    }
  }
  for (uint32_t i = (vecs << 4) + sub; i < n; i += G) dst[i] = __ldg(src + i);
}

// LZ77 match copy by a group: dst[i] = dst[i - off], i in [0, len), sequential-copy
// semantics.  Earlier writes of the group are visible (caller synchronised).
template <int G>
__device__ __forceinline__ void group_copy_match(uint8_t* dst, uint32_t off, uint32_t len, uint32_t sub,
                                                 unsigned gmask)
{
  const uint8_t* src = dst - off;
  if (off >= 16u * G && len >= 16u * G) {
    // 16 bytes per lane and step; a step only reads bytes stored by earlier steps
    const uint32_t head = (uint32_t)((16 - (reinterpret_cast<uintptr_t>(dst) & 15)) & 15);
    for (uint32_t i = sub; i < head; i += G) dst[i] = src[i]; // head < 16 <= off: no overlap inside
    __syncwarp(gmask);
    const uint32_t n = len - head;
    const uint32_t vecs = n >> 4;
    uint8_t* d0 = dst + head;
    const uintptr_t sa = reinterpret_cast<uintptr_t>(src + head);
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(sa & ~(uintptr_t)3);
    const uint32_t sh = (uint32_t)(sa & 3) * 8;
    uint4* dv = reinterpret_cast<uint4*>(d0);
    const bool overlap = off < len;
    for (uint32_t base = 0; base < vecs; base += G) {
      const uint32_t v = base + sub;
      if (v < vecs) {
        const uint32_t* s = sw + 4 * v;
        const uint32_t a = s[0], b = s[1], c = s[2], d = s[3];
        uint4 r;
        if (sh == 0) {
          r = make_uint4(a, b, c, d);
        } else {
          const uint32_t e = s[4]; // bytes below d0 + 16 v + 16 - off + 3 < this step's first store
          r.x = __funnelshift_r(a, b, sh);
          r.y = __funnelshift_r(b, c, sh);
          r.z = __funnelshift_r(c, d, sh);
          r.w = __funnelshift_r(d, e, sh);
        }
        dv[v] = r;
      }
      if (overlap) __syncwarp(gmask);
    }
    for (uint32_t i = (vecs << 4) + sub; i < n; i += G) d0[i] = (d0 - off)[i]; // tail < 16 <= off
    return;
  }
  if (off >= len) {
    for (uint32_t i = sub; i < len; i += G) dst[i] = src[i];
  } else if (off >= (uint32_t)G) {
    for (uint32_t base = 0; base < len; base += G) {
      const uint32_t i = base + sub;
      if (i < len) dst[i] = src[i];
      __syncwarp(gmask);
    }
  } else {
    // period `off` < G, all pattern bytes precede dst
    uint32_t idx = sub % off;
    const uint32_t step = (uint32_t)G % off;
    for (uint32_t i = sub; i < len; i += G) {
      dst[i] = src[idx];
      idx += step;
      if (idx >= off) idx -= off;
    }
  }
}

} // namespace b200comp

// ---------------------------------------------------------------- round executor
//
// The v3 decoders split a chunk's work into rounds of up to 32 elements.  The token
// stream of a round is parsed first (a short dependent chain that only touches the
// compressed bytes); after that every lane holds one element
//     output [pos, lit_end)   <- literals  comp[p + lit_delta]
//     output [lit_end, end)   <- match     out[p - off]
// and the warp produces the round's output span strictly in address order, 32
// consecutive bytes per step: lane t finds the element that owns byte b + t, loads
// its source byte and stores it.  A match byte is ready when its source lies below
// the step's first byte (everything below is complete); the step is cut in front of
// the first byte that is not.  Stores of a step are 32 consecutive bytes.
namespace b200comp {

__device__ __forceinline__ uint32_t add_sat_u32(uint32_t a, uint32_t b)
{
  const uint32_t s = a + b;
  return s < a ? 0xFFFFFFFFu : s;
}

// exclusive prefix sum over the warp, saturating at 2^32 - 1
__device__ __forceinline__ uint32_t warp_excl_sum_sat(uint32_t v, uint32_t& total)
{
  uint32_t incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(kFull, incl, d);
    if (lane_id() >= d) incl = add_sat_u32(incl, o);
  }
  total = __shfl_sync(kFull, incl, 31);
  const uint32_t prev = __shfl_up_sync(kFull, incl, 1);
  return lane_id() == 0 ? 0u : prev;
}

constexpr uint32_t kLongRun = 96; // a segment this long leaves the byte-per-lane loop

__device__ __forceinline__ void execute_round(uint8_t* out, const uint8_t* comp, uint32_t begin, uint32_t end,
                                              uint32_t pos, uint32_t lit_end, uint32_t seq_end,
                                              uint32_t lit_delta, uint32_t off)
{
  const uint32_t lane = (uint32_t)lane_id();
  const unsigned le_mask = 0xFFFFFFFFu >> (31u - lane); // bits 0..lane
  uint32_t b = begin, carry = 0;
  // Round whose matches all reach at least 64 bytes back and that has no long runs: an
  // iteration produces 64 consecutive bytes, so every match byte it needs lies below the
  // iteration's first byte, i.e. was stored by an earlier iteration (made visible by the
  // __syncwarp() that ends each one) or an earlier round.  Two 32-byte steps are in
  // flight per iteration and their loads overlap.
  const bool far = seq_end == lit_end || off >= 64u;
  const bool plain = far && (seq_end - pos) < kLongRun;
  if (__all_sync(kFull, plain)) {
    while (b < end) {
      const uint32_t rel = pos - b;
      const bool mine = seq_end > pos;
      const unsigned s0 = __reduce_or_sync(kFull, (rel < 32u && mine) ? (1u << rel) : 0u);
      const unsigned s1 = __reduce_or_sync(kFull, (rel - 32u < 32u && mine) ? (1u << (rel - 32u)) : 0u);
      const uint32_t c1 = carry + (uint32_t)__popc(s0);
      const uint32_t own0 = carry + (uint32_t)__popc(s0 & le_mask) - 1u;
      const uint32_t own1 = c1 + (uint32_t)__popc(s1 & le_mask) - 1u;
      const uint32_t le0 = __shfl_sync(kFull, lit_end, (int)own0), le1 = __shfl_sync(kFull, lit_end, (int)(own1 & 31u));
      const uint32_t of0 = __shfl_sync(kFull, off, (int)own0), of1 = __shfl_sync(kFull, off, (int)(own1 & 31u));
      const uint32_t dl0 = __shfl_sync(kFull, lit_delta, (int)own0), dl1 = __shfl_sync(kFull, lit_delta, (int)(own1 & 31u));
      const uint32_t p0 = b + lane, p1 = p0 + 32u;
      const uint8_t* a0 = p0 < le0 ? comp + (uint32_t)(p0 + dl0) : out + (p0 - of0);
      const uint8_t* a1 = p1 < le1 ? comp + (uint32_t)(p1 + dl1) : out + (p1 - of1);
      uint8_t v0 = 0, v1 = 0;
      if (p0 < end) v0 = *a0;
      if (p1 < end) v1 = *a1;
      if (p0 < end) out[p0] = v0;
      if (p1 < end) out[p1] = v1;
      carry = c1 + (uint32_t)__popc(s1);
      b += 64u;
      __syncwarp();
    }
    return;
  }
  while (b < end) {
    const uint32_t rel = pos - b; // elements that started below b wrap to a large value
    const unsigned starts = __reduce_or_sync(kFull, (rel < 32u && seq_end > pos) ? (1u << rel) : 0u);
    const uint32_t own = carry + (uint32_t)__popc(starts & le_mask) - 1u;
    const uint32_t o_lit_end = __shfl_sync(kFull, lit_end, (int)own);
    const uint32_t o_off = __shfl_sync(kFull, off, (int)own);
    const uint32_t o_delta = __shfl_sync(kFull, lit_delta, (int)own);
    if ((starts & ~1u) == 0u) {
      // one element owns the whole step: long literal / match runs move as vectors
      const uint32_t o_seq_end = __shfl_sync(kFull, seq_end, (int)own);
      const bool in_lit = b < o_lit_end;
      const uint32_t seg_end = in_lit ? o_lit_end : o_seq_end;
      if (seg_end - b >= kLongRun) {
        if (in_lit)
          group_copy_ro<32>(out + b, comp + (uint32_t)(b + o_delta), seg_end - b, lane);
        else
          group_copy_match<32>(out + b, o_off, seg_end - b, lane, kFull);
        b = seg_end;
        carry += (starts & 1u);
        __syncwarp();
        continue;
      }
    }
    const uint32_t p = b + lane;
    const bool active = p < end;
    const bool is_lit = p < o_lit_end;
    uint32_t q = p - o_off;
    bool ready = is_lit || q < b;
    if (active && !ready) {
      // source inside this step: fold it back by whole periods if that stays inside
      // the match's own source window [lit_end - off, ...)
      const uint32_t n = (q - b) / o_off + 1u;
      const uint32_t q2 = q - n * o_off;
      if (q2 + o_off >= o_lit_end) {
        q = q2;
        ready = true;
      }
    }
    const unsigned blocked = __ballot_sync(kFull, active && !ready);
    const uint32_t limit = blocked ? (uint32_t)__ffs((int)blocked) - 1u : 32u;
    if (active && lane < limit) {
      const uint8_t* src = is_lit ? comp + (uint32_t)(p + o_delta) : out + q;
      out[p] = *src;
    }
    const uint32_t adv = min(limit, end - b);
    carry += (uint32_t)__popc(starts & (adv >= 32u ? 0xFFFFFFFFu : ((1u << adv) - 1u)));
    b += adv;
    __syncwarp();
  }
}


// ---- round executor with a start bitmap (the fast path of the window decoders).
// Condition: the round produces at most kOwnCap bytes; all literals of the round lie in the shared-memory
// window (literal byte p of a sequence is window byte p + win_delta).
// own[] (shared memory, kOwnWords words per warp) gets a bitmap of the output bytes where a sequence starts, one
// word per 32 bytes, and per word the number of starts in front of it: the owner of byte 32 w + l is then
// base[w] + popc(bits[w] & (2 << l) - 1) - 1, two broadcast loads and three ALU instructions per 32 bytes.
// A step produces 64 bytes.  A match byte whose source lies inside the match itself (distance shorter than the
// bytes already produced) is folded back by whole periods to a byte in front of the match; a source that still
// lies inside the step (a neighbour sequence produced by other lanes of the same step) is taken from a 64-byte
// stage in shared memory once the lane that produces it has done so - passes until the step is complete, one
// for steps without such sources.
// Returns false (nothing written) when the round does not qualify; the caller falls back to execute_round().
constexpr uint32_t kOwnCap = 1024;      // bytes of output per round on this path
constexpr uint32_t kOwnWords = 64 + 16; // shared-memory words per warp: 32 bitmap words, 32 bases, 64 stage bytes

__device__ __forceinline__ uint32_t lds_u8_shared(uint32_t sa)
{
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(sa));
  return v;
}

__device__ __forceinline__ bool execute_round_owned(uint8_t* out, const uint8_t* win, uint32_t* own, uint32_t begin,
                                                    uint32_t end, uint32_t pos, uint32_t lit_end, uint32_t seq_end,
                                                    uint32_t win_delta, uint32_t off)
{
  const uint32_t lane = (uint32_t)lane_id();
  const uint32_t span = end - begin;
  if (span > kOwnCap) return false;
  const unsigned le_mask = 0xFFFFFFFFu >> (31u - lane); // bits 0..lane
  own[lane] = 0u;
  __syncwarp();
  const uint32_t rel = pos - begin;
  if (seq_end > pos) atomicOr(&own[rel >> 5], 1u << (rel & 31u));
  __syncwarp();
  {
    const uint32_t cnt = (uint32_t)__popc(own[lane]);
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(kFull, incl, d);
      if ((int)lane >= d) incl += t;
    }
    own[32u + lane] = incl - cnt - 1u; // starts in front of the word, minus one: owner = this + starts up to the byte
  }
  __syncwarp();
  // every match of the round reaches at least 64 bytes back: a 64-byte step never reads what it writes
  const bool far = seq_end == lit_end || off >= 64u;
  if (__all_sync(kFull, far)) {
    const uint32_t win_sa = (uint32_t)__cvta_generic_to_shared(win) + win_delta; // + p = the literal's window byte
    for (uint32_t b = 0; b < span; b += 64u) {
      const uint32_t w = b >> 5;
      const uint32_t o0 = own[32u + w] + (uint32_t)__popc(own[w] & le_mask);
      const uint32_t o1 = own[33u + w] + (uint32_t)__popc(own[w + 1u] & le_mask);
      const uint32_t p0 = begin + b + lane, p1 = p0 + 32u;
      const bool in0 = p0 < end, in1 = p1 < end;
      const uint32_t le0 = __shfl_sync(kFull, lit_end, (int)(o0 & 31u)), le1 = __shfl_sync(kFull, lit_end, (int)(o1 & 31u));
      const uint32_t of0 = __shfl_sync(kFull, off, (int)(o0 & 31u)), of1 = __shfl_sync(kFull, off, (int)(o1 & 31u));
      const uint32_t wd0 = __shfl_sync(kFull, win_sa, (int)(o0 & 31u)), wd1 = __shfl_sync(kFull, win_sa, (int)(o1 & 31u));
      const bool lit0 = p0 < le0, lit1 = p1 < le1;
      uint32_t v0 = 0, v1 = 0;
      if (in0 && lit0) v0 = lds_u8_shared(wd0 + p0);
      if (in1 && lit1) v1 = lds_u8_shared(wd1 + p1);
      if (in0 && !lit0) v0 = out[p0 - of0];
      if (in1 && !lit1) v1 = out[p1 - of1];
      if (in0) out[p0] = (uint8_t)v0;
      if (in1) out[p1] = (uint8_t)v1;
      __syncwarp();
    }
    return true;
  }
  uint8_t* stage = reinterpret_cast<uint8_t*>(own + 64);
  const uint32_t win_sa = (uint32_t)__cvta_generic_to_shared(win) + win_delta; // + p = the literal's window byte
  // x / off for x < kOwnCap by multiplication (exact: x * off < 2^32 whenever the quotient is needed, x >= off)
  const uint32_t magic = off ? 0xFFFFFFFFu / off + 1u : 0u;
  for (uint32_t b = 0; b < span; b += 64u) {
    const uint32_t w = b >> 5;
    const uint32_t o0 = own[32u + w] + (uint32_t)__popc(own[w] & le_mask);
    const uint32_t o1 = own[33u + w] + (uint32_t)__popc(own[w + 1u] & le_mask);
    const uint32_t b_abs = begin + b;
    const uint32_t p0 = b_abs + lane, p1 = p0 + 32u;
    const bool in0 = p0 < end, in1 = p1 < end;
    const uint32_t le0 = __shfl_sync(kFull, lit_end, (int)(o0 & 31u)), le1 = __shfl_sync(kFull, lit_end, (int)(o1 & 31u));
    const uint32_t of0 = __shfl_sync(kFull, off, (int)(o0 & 31u)), of1 = __shfl_sync(kFull, off, (int)(o1 & 31u));
    const uint32_t wd0 = __shfl_sync(kFull, win_sa, (int)(o0 & 31u)), wd1 = __shfl_sync(kFull, win_sa, (int)(o1 & 31u));
    const uint32_t mg0 = __shfl_sync(kFull, magic, (int)(o0 & 31u)), mg1 = __shfl_sync(kFull, magic, (int)(o1 & 31u));
    const bool lit0 = p0 < le0, lit1 = p1 < le1;
    // source of a match byte: p - off, folded in front of the match when the match overlaps itself
    const uint32_t x0 = p0 - le0, x1 = p1 - le1;
    uint32_t s0 = p0 - of0, s1 = p1 - of1;
    if (!lit0 && x0 >= of0) s0 = le0 - of0 + (x0 - __umulhi(x0, mg0) * of0);
    if (!lit1 && x1 >= of1) s1 = le1 - of1 + (x1 - __umulhi(x1, mg1) * of1);
    bool pend0 = in0 && !lit0 && s0 >= b_abs, pend1 = in1 && !lit1 && s1 >= b_abs;
    uint32_t v0 = 0, v1 = 0;
    if (in0 && lit0) v0 = lds_u8_shared(wd0 + p0);
    if (in1 && lit1) v1 = lds_u8_shared(wd1 + p1);
    if (in0 && !lit0 && !pend0) v0 = out[s0];
    if (in1 && !lit1 && !pend1) v1 = out[s1];
    if (in0 && !pend0) out[p0] = (uint8_t)v0;
    if (in1 && !pend1) out[p1] = (uint8_t)v1;
    unsigned wait0 = __ballot_sync(kFull, pend0), wait1 = __ballot_sync(kFull, pend1);
    if (wait0 | wait1) {
      // sources inside the step: through the stage, in as many passes as the chain of neighbours is long
      if (!pend0) stage[lane] = (uint8_t)v0;
      if (!pend1) stage[32u + lane] = (uint8_t)v1;
      const uint32_t r0 = s0 - b_abs, r1 = s1 - b_abs; // < 64 when pending
      while (wait0 | wait1) {
        __syncwarp();
        const bool go0 = pend0 && !(((r0 < 32u ? wait0 : wait1) >> (r0 & 31u)) & 1u);
        const bool go1 = pend1 && !(((r1 < 32u ? wait0 : wait1) >> (r1 & 31u)) & 1u);
        if (go0) v0 = stage[r0];
        if (go1) v1 = stage[r1];
        __syncwarp();
        if (go0) {
          stage[lane] = (uint8_t)v0;
          out[p0] = (uint8_t)v0;
          pend0 = false;
        }
        if (go1) {
          stage[32u + lane] = (uint8_t)v1;
          out[p1] = (uint8_t)v1;
          pend1 = false;
        }
        const unsigned n0 = __ballot_sync(kFull, pend0), n1 = __ballot_sync(kFull, pend1);
        if (n0 == wait0 && n1 == wait1) break; // cannot happen for a validated round (a source lies below its byte)
        wait0 = n0, wait1 = n1;
      }
    }
    __syncwarp();
  }
  return true;
}

} // namespace b200comp
