// b200comp — LZ4 block encoder for chunks <= 64 KiB: one CTA ("team" of W warps) per chunk.
//
// Replaces the reference's one-warp-per-chunk greedy encoder (src/LZ4Kernels.hiph:793-969,
// launched by src/lowlevel/LZ4CompressionKernels.hip:73-94), whose per-window chain
// "key -> table probe -> verify -> extend" is serial per chunk.  The output is a plain LZ4 block
// (decodable by the reference decoder and by liblz4) but not the reference's byte stream; the
// policy is stated sequentially in oracle/lz4_team_model.c, which the tests compare against byte
// for byte.
//
// Data movement: the chunk is staged in shared memory by one bulk copy (cp.async.bulk +
// mbarrier), the uint16 hash table lives in shared memory, neither touches HBM again.  The chunk
// is cut into segments of 1024 positions, dealt round-robin to the warps.  Per segment a warp
//   1. hashes its positions (parallel across warps),
//   2. takes the "turn" (an mbarrier chain in segment order) and probes + inserts its hashes into
//      the table, 32 positions per step: this is the only part serialised per chunk, ~14 warp
//      instructions per 32 positions, and it yields d[p] = distance to the most recent earlier
//      position with the same hash, exactly as a sequential encoder that inserts every position,
//   3. parses the segment greedily from d[] (verify 32 candidates at once, extend the first hit
//      128 bytes per step, reuse the verified mask behind the match), staging the sequences in
//      the dead part of its own d[] buffer,
//   4. waits for its predecessor's output cursor (second mbarrier chain), writes the staged bytes
//      to global memory with 16-byte stores and passes the cursor on.
// A match that stops at a segment end stays open and is merged with the next segment's first
// match when that starts at its first byte with the same offset (runs / periodic data collapse
// to one sequence, as in the reference).
#pragma once

#include "lz_team.cuh"

namespace b200comp {
namespace lz4 {

using team::kSeg;

template <int CAP, int HL, int W>
struct TeamLayout {
  static constexpr uint32_t kChunkBuf = CAP + 48; // 16 bytes of phase slack + 32 bytes of over-read pad
  static constexpr uint32_t kTableOff = kChunkBuf;
  static constexpr uint32_t kDbufOff = kTableOff + (2u << HL);
  static constexpr uint32_t kCtrlOff = kDbufOff + W * kSeg * 2;
  static constexpr uint32_t kBytes = kCtrlOff + 16 + 48 * W; // load barrier, turn barriers, publish barriers, states
};

// hand-over between consecutive segments: output cursor + what is pending in front of it
struct TeamState {
  uint32_t out_base; // bytes of the chunk's stream written so far
  uint32_t flags;    // bit 0: an open match is pending; bits 16..31: its offset
  uint32_t value;    // open: match length so far; else: first literal not yet written
  uint32_t tok;      // open: position of its token (bits 0..19) and literal nibble (bits 20..23)
};

template <int CAP, int HL, int W>
__global__ void __launch_bounds__(W * 32, (CAP > 16384 ? 2 : (CAP > 4096 ? 3 : 6)))
compress_team_kernel(const void* const* in_ptrs, const size_t* in_bytes, void* const* out_ptrs, size_t* out_bytes,
                     size_t batch)
{
  using L = TeamLayout<CAP, HL, W>;
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* s_chunk = smem;
  const uint32_t* s_chunk32 = reinterpret_cast<const uint32_t*>(smem);
  uint16_t* s_table = reinterpret_cast<uint16_t*>(smem + L::kTableOff);
  uint16_t* s_dbuf = reinterpret_cast<uint16_t*>(smem + L::kDbufOff);
  uint64_t* load_bar = reinterpret_cast<uint64_t*>(smem + L::kCtrlOff);
  uint64_t* turn_bar = load_bar + 2;          // [W], one 16-byte slot each
  uint64_t* pub_bar = turn_bar + 2 * W;       // [W]
  uint4* s_state = reinterpret_cast<uint4*>(pub_bar + 2 * W); // [W]

  const uint32_t lane = (uint32_t)lane_id();
  const uint32_t warp = threadIdx.x >> 5;
  const unsigned lt_mask = lanemask_lt();

  if (threadIdx.x == 0) {
    team::mbar_init(load_bar, 1);
    for (int w = 0; w < W; ++w) {
      team::mbar_init(turn_bar + 2 * w, 1);
      team::mbar_init(pub_bar + 2 * w, 1);
    }
    team::mbar_init_fence();
  }
  uint32_t load_phase = 0, turn_phase = 0, pub_phase = 0;

  for (size_t chunk = blockIdx.x; chunk < batch; chunk += gridDim.x) {
    __syncthreads(); // the previous chunk is finished with shared memory (and the barriers are initialised)
    const uint8_t* src = static_cast<const uint8_t*>(in_ptrs[chunk]);
    uint8_t* gout = static_cast<uint8_t*>(out_ptrs[chunk]);
    const size_t n64 = in_bytes[chunk];
    if (n64 == 0 || n64 > (size_t)CAP) { // empty chunk -> empty stream; larger than the caller's declared maximum -> 0
      if (threadIdx.x == 0) out_bytes[chunk] = 0;
      continue;
    }
    const uint32_t n = (uint32_t)n64;
    const uint32_t mis = team::stage_chunk(s_chunk, src, n, load_bar);
    {
      uint4* t128 = reinterpret_cast<uint4*>(s_table);
      const uint4 ones = make_uint4(~0u, ~0u, ~0u, ~0u);
      for (uint32_t i = threadIdx.x; i < (2u << HL) / 16u; i += W * 32) t128[i] = ones;
    }
    team::mbar_wait(load_bar, load_phase);
    load_phase ^= 1u;
    __syncthreads();

    const uint32_t npos = n >= 12u ? n - 11u : 0u;  // positions that may start a match: p <= n - 12
    const uint32_t mlimit = n >= 5u ? n - 5u : 0u;  // matches end at or before n - 5
    const uint32_t nseg = (n + kSeg - 1u) / kSeg;
    uint16_t* db = s_dbuf + warp * kSeg;
    uint8_t* stage = reinterpret_cast<uint8_t*>(db);
    const uint32_t* stage32 = reinterpret_cast<const uint32_t*>(db);
    const uint32_t nextw = (warp + 1u) % W;
    const uint32_t chunk_sa = team::smem_addr(s_chunk) + mis; // shared address of chunk byte 0
    const uint32_t db_sa = team::smem_addr(db);

    for (uint32_t s = warp; s < nseg; s += W) {
      const uint32_t S0 = s * kSeg;
      const uint32_t S1 = min(S0 + kSeg, n);
      const uint32_t E = min(mlimit, S1);
      const uint32_t pend = min(S1, npos);
      const uint32_t cnt = pend > S0 ? pend - S0 : 0u;

      // ---- 1. hash every position; positions that repeat a hash of a lower lane of their aligned
      //         32-window are settled here (candidate = nearest such lane, record = 0x8000 | distance),
      //         the others keep their hash for the table (record = hash)
      {
        uint32_t w0 = 0;
#pragma unroll 2
        for (; w0 + 32 <= cnt; w0 += 32) {
          const uint32_t key = team::lds_u32_at(s_chunk32, mis + S0 + w0 + lane);
          const uint32_t h = (key * 2654435761u) >> (32 - HL);
          const unsigned lower = __match_any_sync(kFull, h) & lt_mask;
          db[w0 + lane] = (uint16_t)(lower ? (0x8000u | (lane - (31u - (uint32_t)__clz((int)lower)))) : h);
        }
        if (w0 < cnt) {
          const uint32_t i = w0 + lane;
          const bool valid = i < cnt;
          const unsigned vm = __ballot_sync(kFull, valid);
          if (valid) {
            const uint32_t key = team::lds_u32_at(s_chunk32, mis + S0 + i);
            const uint32_t h = (key * 2654435761u) >> (32 - HL);
            const unsigned lower = __match_any_sync(vm, h) & lt_mask;
            db[i] = (uint16_t)(lower ? (0x8000u | (lane - (31u - (uint32_t)__clz((int)lower)))) : h);
          }
        }
      }
      __syncwarp();

      // ---- 2. table probe + insert, in segment order across the team (the turn chain).  The lanes that
      //         touch the table in one step have distinct hashes, so the step is order-free inside.
      if (s > 0) {
        team::mbar_wait(turn_bar + 2 * warp, turn_phase);
        turn_phase ^= 1u;
      }
      for (uint32_t w0 = 0; w0 < cnt; w0 += 128) {
        uint32_t rec[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t i = w0 + 32 * u + lane;
          rec[u] = i < cnt ? (uint32_t)db[i] : 0x8000u;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (!(rec[u] & 0x8000u)) {
            const uint32_t pp = S0 + w0 + 32 * u + lane;
            const uint32_t t = s_table[rec[u]];
            s_table[rec[u]] = (uint16_t)pp;
            rec[u] = t < pp ? pp - t : 0u;
          } else {
            rec[u] &= 0x7FFFu;
          }
          __syncwarp();
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint32_t i = w0 + 32 * u + lane;
          if (i < cnt) db[i] = (uint16_t)rec[u];
        }
      }
      __syncwarp();
      if (s + 1 < nseg) {
        if (lane == 0) team::mbar_arrive(turn_bar + 2 * nextw);
      }

      // ---- 3. parse; sequences are collected one per lane and staged 32 at a time over the consumed
      //         part of db[]
      uint32_t sp = 0;               // staged bytes
      uint32_t lit = S0;             // first literal not yet covered by a sequence
      uint32_t nb = 0;               // sequences collected in r_* (lane k holds sequence k)
      uint32_t r_lit = 0, r_q = 0, r_mend = 0, r_d = 0;
      bool have_first = false;       // the first sequence's token + literals are written at flush time
      uint32_t f_q = 0, f_mlen = 0, f_d = 0, f_skip = 0;
      bool l_open = false, l_first = false; // the last match stops at S1
      uint32_t l_tok_sp = 0, l_hi = 0, l_mlen = 0, l_d = 0;

      auto emit_batch = [&]() {
        const bool act = lane < nb;
        const bool isf = !have_first && lane == 0;
        if (!have_first) {
          have_first = true;
          f_q = __shfl_sync(kFull, r_q, 0);
          f_mlen = __shfl_sync(kFull, r_mend, 0) - f_q;
          f_d = __shfl_sync(kFull, r_d, 0);
        }
        const uint32_t nlit = r_q - r_lit, mlen = r_mend - r_q;
        const bool open = act && r_mend == S1;
        const uint32_t hi = nlit >= 15u ? 15u : nlit;
        uint32_t size = 0;
        if (act) {
          if (!isf) size = 1u + nlit + (nlit >= 15u ? (nlit - 15u) / 255u + 1u : 0u);
          size += 2u + ((!open && mlen >= 19u) ? (mlen - 19u) / 255u + 1u : 0u);
        }
        uint32_t total;
        const uint32_t off = warp_excl_sum(size, total);
        uint32_t o = sp + off, lit_dst = 0;
        if (act) {
          if (!isf) {
            stage[o++] = (uint8_t)((hi << 4) | (open ? 0u : (mlen - 4u >= 15u ? 15u : mlen - 4u)));
            if (nlit >= 15u) {
              uint32_t x = nlit - 15u;
              for (; x >= 255u; x -= 255u) stage[o++] = 0xff;
              stage[o++] = (uint8_t)x;
            }
            lit_dst = o;
            if (nlit <= 16u)
              for (uint32_t i = 0; i < nlit; ++i) stage[o + i] = s_chunk[mis + r_lit + i];
            o += nlit;
          }
          stage[o] = (uint8_t)(r_d & 0xffu);
          stage[o + 1] = (uint8_t)(r_d >> 8);
          o += 2u;
          if (!open && mlen >= 19u) {
            uint32_t x = mlen - 19u;
            for (; x >= 255u; x -= 255u) stage[o++] = 0xff;
            stage[o++] = (uint8_t)x;
          }
        }
        // literal runs above 16 bytes: the whole warp copies them, one run at a time
        unsigned longm = __ballot_sync(kFull, act && !isf && nlit > 16u);
        while (longm) {
          const int l = __ffs((int)longm) - 1;
          longm &= longm - 1u;
          const uint32_t src = __shfl_sync(kFull, r_lit, l), cntl = __shfl_sync(kFull, nlit, l),
                         dst = __shfl_sync(kFull, lit_dst, l);
          for (uint32_t i = lane; i < cntl; i += 32) stage[dst + i] = s_chunk[mis + src + i];
        }
        if (sp == 0u) f_skip = __shfl_sync(kFull, size, 0); // bytes the first sequence of the segment staged
        const unsigned openm = __ballot_sync(kFull, open);
        if (openm) {
          const int l = (int)nb - 1; // only the last sequence of a segment can stop at S1
          l_open = true;
          l_first = sp == 0u && l == 0;
          l_tok_sp = sp + __shfl_sync(kFull, off, l);
          l_hi = __shfl_sync(kFull, hi, l);
          l_mlen = __shfl_sync(kFull, mlen, l);
          l_d = __shfl_sync(kFull, r_d, l);
        }
        sp += total;
        nb = 0;
        __syncwarp();
      };

      uint32_t c = S0;
      while (c < pend) {
        const uint32_t p = c + lane;
        uint32_t d = 0;
        if (p < pend) d = team::lds_u16(db_sa + 2u * (p - S0));
        bool ok = false;
        if (d != 0u && p + 4u <= E)
          ok = team::lds_u32_unaligned(chunk_sa + p) == team::lds_u32_unaligned(chunk_sa + p - d);
        const unsigned vmask = __ballot_sync(kFull, ok);
        if (!vmask) {
          c += 32;
          continue;
        }
        // a run of consecutive positions that match at the same distance is one match: the run itself
        // proves [q, run end + 3) equal, the bytes behind it are compared 32 per step.  Every lane
        // prepares the first byte to compare for a match that would start at it.
        const uint32_t dprev = __shfl_up_sync(kFull, d, 1);
        const bool cont = ok && lane > 0u && dprev == d && ((vmask >> (lane - 1u)) & 1u);
        const unsigned cmask = __ballot_sync(kFull, cont);
        const unsigned above = ~cmask & (0xFFFFFFFEu << lane);
        const uint32_t cmp0 = c + 3u + (above ? (uint32_t)__ffs((int)above) - 1u : 32u);
        // greedy walk over the window: first verified position at or behind the cursor, jump to its end
        unsigned m = vmask;
        for (;;) {
          const int ql = __ffs((int)m) - 1;
          const uint32_t dq = __shfl_sync(kFull, d, ql);
          uint32_t a0 = __shfl_sync(kFull, cmp0, ql);
          uint32_t mq;
          for (;; a0 += 32u) {
            const uint32_t a = a0 + lane;
            bool neq = true;
            if (a < E) neq = team::lds_u8(chunk_sa + a) != team::lds_u8(chunk_sa + a - dq);
            const unsigned diff = __ballot_sync(kFull, neq);
            if (diff) {
              mq = a0 + (uint32_t)__ffs((int)diff) - 1u;
              break;
            }
          }
          if (lane == nb) {
            r_lit = lit;
            r_q = c + (uint32_t)ql;
            r_mend = mq;
            r_d = dq;
          }
          ++nb;
          lit = mq;
          if (nb == 32u) emit_batch();
          const uint32_t rel = mq - c;
          if (rel >= 32u) break;
          m = vmask & (0xFFFFFFFFu << rel);
          if (!m) break;
        }
        c = max(lit, c + 32u);
      }
      if (nb) emit_batch();
      __syncwarp();

      // ---- 4. flush behind the predecessor's cursor
      TeamState st;
      if (s > 0) {
        team::mbar_wait(pub_bar + 2 * warp, pub_phase);
        pub_phase ^= 1u;
        const uint4 v = s_state[warp];
        st.out_base = v.x;
        st.flags = v.y;
        st.value = v.z;
        st.tok = v.w;
      } else {
        st.out_base = 0;
        st.flags = 0;
        st.value = 0;
        st.tok = 0;
      }
      uint32_t ob = st.out_base;
      TeamState ns; // state handed to the next segment
      bool whole = false; // the segment is one continuation of the pending match, which stays open
      bool o_open = (st.flags & 1u) != 0u;
      uint32_t lit_from = st.value;
      uint32_t stage_from = 0;
      bool header_done = false; // the first sequence needs no token + literals (merged into the pending match)
      if (o_open) {
        const uint32_t o_d = st.flags >> 16;
        uint32_t total = st.value;
        bool merge = have_first && f_q == S0;
        if (merge && f_d != o_d) {
          // the pending match's own offset must reproduce this whole match
          for (uint32_t k = 0; k < f_mlen && merge; k += 128u) {
            const uint32_t a = S0 + k + 4u * lane;
            uint32_t x = 0;
            if (a < S0 + f_mlen) {
              x = team::lds_u32_at(s_chunk32, mis + a) ^ team::lds_u32_at(s_chunk32, mis + a - o_d);
              const uint32_t left = S0 + f_mlen - a;
              if (left < 4u) x &= (1u << (8u * left)) - 1u;
            }
            if (__any_sync(kFull, x != 0u)) merge = false;
          }
        }
        if (merge) total += f_mlen;
        if (merge && l_open && l_first) {
          whole = true;
          ns.out_base = ob;
          ns.flags = st.flags;
          ns.value = total;
          ns.tok = st.tok;
        } else {
          // close the pending match with its final length
          const uint32_t m = total - 4u;
          if (lane == 0) gout[st.tok & 0xFFFFFu] = (uint8_t)(((st.tok >> 20) << 4) | (m >= 15u ? 15u : m));
          if (m >= 15u) ob += team::put_lsic(gout + ob, m - 15u);
          if (merge) {
            header_done = true;
            stage_from = f_skip;
          }
          lit_from = S0;
        }
      }
      if (!whole) {
        if (!have_first) {
          ns.out_base = ob;
          ns.flags = 0;
          ns.value = lit_from;
          ns.tok = 0;
        } else {
          uint32_t first_tok = 0, first_hi = 0;
          if (!header_done) {
            const uint32_t nlit = f_q - lit_from;
            first_hi = nlit >= 15u ? 15u : nlit;
            first_tok = ob;
            if (lane == 0)
              gout[ob] = (uint8_t)((first_hi << 4) | ((l_open && l_first) ? 0u : (f_mlen - 4u >= 15u ? 15u : f_mlen - 4u)));
            ob += 1u;
            if (nlit >= 15u) ob += team::put_lsic(gout + ob, nlit - 15u);
            team::copy_s2g(gout + ob, s_chunk32, mis + lit_from, nlit);
            ob += nlit;
          }
          const uint32_t stage_base = ob - stage_from; // stream position of staged byte 0
          team::copy_s2g(gout + ob, stage32, stage_from, sp - stage_from);
          ob += sp - stage_from;
          ns.out_base = ob;
          if (l_open) {
            ns.flags = 1u | (l_d << 16);
            ns.value = l_mlen;
            ns.tok = l_first ? (first_tok | (first_hi << 20)) : ((stage_base + l_tok_sp) | (l_hi << 20));
          } else {
            ns.flags = 0;
            ns.value = lit;
            ns.tok = 0;
          }
        }
      }
      __syncwarp();
      if (s + 1 < nseg) {
        if (lane == 0) {
          s_state[nextw] = make_uint4(ns.out_base, ns.flags, ns.value, ns.tok);
          team::mbar_arrive(pub_bar + 2 * nextw);
        }
      } else {
        // last segment: the final sequence is literals only
        uint32_t o2 = ns.out_base;
        uint32_t from = ns.value;
        if (ns.flags & 1u) { // unreachable (matches end at n - 5 < S1 here); closed for safety
          const uint32_t m = ns.value - 4u;
          if (lane == 0) gout[ns.tok & 0xFFFFFu] = (uint8_t)(((ns.tok >> 20) << 4) | (m >= 15u ? 15u : m));
          if (m >= 15u) o2 += team::put_lsic(gout + o2, m - 15u);
          from = S1;
        }
        const uint32_t nlit = n - from;
        if (lane == 0) gout[o2] = (uint8_t)((nlit >= 15u ? 15u : nlit) << 4);
        o2 += 1u;
        if (nlit >= 15u) o2 += team::put_lsic(gout + o2, nlit - 15u);
        team::copy_s2g(gout + o2, s_chunk32, mis + from, nlit);
        o2 += nlit;
        if (lane == 0) out_bytes[chunk] = o2;
      }
      __syncwarp();
    }
  }
}

} // namespace lz4
} // namespace b200comp
