// b200comp — v4 LZ4 / Snappy decoder: parsing decoupled from copying.
//
// A CTA is one PARSE warp plus kExecWarps EXEC warps.  Lane i of the parse warp walks
// the token stream of the chunk that exec warp i is producing, so the serial part of
// the format (token -> lengths -> next token) runs SIMD across chunks: one warp
// instruction advances up to kExecWarps streams.  Parsed elements go, 32 at a time
// (a "round"), into a small shared-memory ring per exec warp; the exec warp turns a
// round into output bytes with the in-order round executor (lz_common.cuh), 32
// consecutive bytes per step.  Validation (bounds, offsets, capacity) is done once,
// by the parse lane, before a round is published: the exec side never sees an
// element it may not run.
//
// Ring protocol (all in shared memory, one producer lane / one consumer warp):
//   head[i] = rounds published, tail[i] = rounds consumed; the producer fills slot
//   head % kRing while head - tail < kRing, fences, then bumps head; the consumer
//   spins on head, copies its element into registers, runs it, then bumps tail.
#pragma once

#include "lz_common.cuh"

namespace b200comp {
namespace pipe {

constexpr int kExecWarps = 15;
constexpr int kThreads = (kExecWarps + 1) * 32;
constexpr int kRing = 2;
constexpr int kCtasPerSm = 3;

enum : uint32_t { kFirst = 1u, kLast = 2u, kError = 4u, kExit = 8u };
enum : int { kToken = 0, kDone = 1, kFail = 2 };

// One element: output [pos, lit_end) <- comp[lp ...], [lit_end, next pos) <- out[. - off]
struct __align__(16) Desc {
  uint32_t pos, lit_end, lp, off;
};

struct __align__(16) RoundMeta {
  uint32_t count, begin, end, flags;
  uint32_t chunk_lo, chunk_hi, result, pad;
};

struct Slot {
  Desc desc[32];
  RoundMeta meta;
};

struct Shared {
  Slot ring[kExecWarps][kRing];
  volatile uint32_t head[kExecWarps];
  volatile uint32_t tail[kExecWarps];
};

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// ------------------------------------------------------------------ exec side
__device__ __forceinline__ void exec_warp(Shared& sh, int w, const void* const* comp_ptrs, void* const* out_ptrs,
                                          size_t* actual, hipcompStatus_t* statuses)
{
  const uint32_t lane = (uint32_t)lane_id();
  const uint8_t* comp = nullptr;
  uint8_t* out = nullptr;
  for (uint32_t r = 0;; ++r) {
    while (sh.head[w] <= r) __nanosleep(20);
    __threadfence_block();
    const Slot& slot = sh.ring[w][r % kRing];
    const RoundMeta m = slot.meta;
    const Desc d = slot.desc[lane];
    if (m.flags & kExit) break;
    const size_t chunk = ((size_t)m.chunk_hi << 32) | m.chunk_lo;
    if (m.flags & kFirst) {
      comp = static_cast<const uint8_t*>(comp_ptrs[chunk]);
      out = static_cast<uint8_t*>(out_ptrs[chunk]);
    }
    const bool have = lane < m.count;
    const uint32_t pos = have ? d.pos : m.end;
    const uint32_t lit_end = have ? d.lit_end : m.end;
    const uint32_t nxt = __shfl_down_sync(kFull, pos, 1);
    const uint32_t seq_end = (lane == 31 || !have) ? m.end : nxt; // lanes past count hold m.end
    execute_round(out, comp, m.begin, m.end, pos, lit_end, seq_end, d.lp - pos, d.off);
    __syncwarp();
    if (lane == 0) {
      if (m.flags & kLast) {
        if (actual) actual[chunk] = m.result;
        if (statuses) statuses[chunk] = (m.flags & kError) ? hipcompErrorCannotDecompress : hipcompSuccess;
      }
      sh.tail[w] = r + 1;
    }
  }
}

// ------------------------------------------------------------------ parse side
//
// Codec contract:
//   struct State;                                   per-stream registers
//   static void open(State&, comp, comp_len, cap);  start a chunk (may already fail)
//   static int  step(State&, Desc&);                kToken: one validated element, state advanced
//                                                   kDone / kFail: nothing produced, state untouched
//   static uint32_t out_pos(const State&);          output bytes produced so far
//   static uint32_t result(const State&, bool ok);  value for device_actual_uncompressed_bytes
template <class Codec>
__device__ __forceinline__ void parse_warp(Shared& sh, const void* const* comp_ptrs, const size_t* comp_bytes,
                                           const size_t* out_caps, void* const* out_ptrs, size_t batch)
{
  const uint32_t i = (uint32_t)lane_id();
  const size_t stride = (size_t)gridDim.x * kExecWarps;
  size_t chunk = (size_t)blockIdx.x * kExecWarps + i;
  typename Codec::State st;
  const uint8_t* out = nullptr;
  bool alive = i < (uint32_t)kExecWarps, open = false, first = false;
  uint32_t r = 0;
  for (;;) {
    const bool can = alive && (r - sh.tail[i < (uint32_t)kExecWarps ? i : 0] < (uint32_t)kRing);
    if (__ballot_sync(kFull, can) == 0u) {
      if (!__any_sync(kFull, alive)) break;
      __nanosleep(40);
      continue;
    }
    if (can) {
      Slot& slot = sh.ring[i][r % kRing];
      if (!open) {
        if (chunk >= batch) {
          slot.meta.flags = kExit;
          __threadfence_block();
          sh.head[i] = r + 1;
          alive = false;
        } else {
          out = static_cast<const uint8_t*>(out_ptrs[chunk]);
          Codec::open(st, static_cast<const uint8_t*>(comp_ptrs[chunk]), (uint32_t)comp_bytes[chunk], out_caps[chunk]);
          open = true;
          first = true;
        }
      }
      if (open) {
        const uint32_t begin = Codec::out_pos(st);
        uint32_t k = 0;
        int rc = kToken;
        while (k < 32u) {
          Desc d;
          rc = Codec::step(st, d);
          if (rc != kToken) break;
          slot.desc[k] = d;
          ++k;
          if (d.off >= 1024u) prefetch_l1(out + (d.lit_end - d.off)); // far source: have it in L1 by exec time
        }
        const bool last = rc != kToken;
        RoundMeta m;
        m.count = k;
        m.begin = begin;
        m.end = Codec::out_pos(st);
        m.flags = (first ? kFirst : 0u) | (last ? kLast : 0u) | (rc == kFail ? kError : 0u);
        m.chunk_lo = (uint32_t)chunk;
        m.chunk_hi = (uint32_t)((unsigned long long)chunk >> 32);
        m.result = Codec::result(st, rc != kFail);
        m.pad = 0;
        slot.meta = m;
        __threadfence_block();
        sh.head[i] = ++r;
        first = false;
        if (last) {
          open = false;
          chunk += stride;
        }
      } else {
        ++r; // exit round published above
      }
    }
  }
}

template <class Codec>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
decompress_pipeline_kernel(const void* const* comp_ptrs, const size_t* comp_bytes, const size_t* out_caps,
                           void* const* out_ptrs, size_t* actual, hipcompStatus_t* statuses, size_t batch)
{
  __shared__ Shared sh;
  if (threadIdx.x < kExecWarps) {
    sh.head[threadIdx.x] = 0;
    sh.tail[threadIdx.x] = 0;
  }
  __syncthreads();
  const int w = threadIdx.x >> 5;
  if (w == kExecWarps)
    parse_warp<Codec>(sh, comp_ptrs, comp_bytes, out_caps, out_ptrs, batch);
  else
    exec_warp(sh, w, comp_ptrs, out_ptrs, actual, statuses);
}

} // namespace pipe
} // namespace b200comp
