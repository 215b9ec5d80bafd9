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
constexpr int kLaneSteps = 4; // 4-byte steps a lane extends its own match before the warp takes over

template <int CAP, int HL, int W>
struct TeamLayout {
  static constexpr uint32_t kFront = 16;          // readable bytes in front of the chunk (phase 1 looks 8 back)
  static constexpr uint32_t kChunkBuf = kFront + CAP + 48; // + 16 bytes of phase slack + 32 bytes of over-read pad
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
  uint8_t* s_chunk = smem + L::kFront;
  const uint32_t* s_chunk32 = reinterpret_cast<const uint32_t*>(smem + L::kFront);
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
    const uint32_t nextw = (warp + 1u) % W;
    const uint32_t chunk_sa = team::smem_addr(s_chunk) + mis; // shared address of chunk byte 0
    const uint32_t db_sa = team::smem_addr(db);

    for (uint32_t s = warp; s < nseg; s += W) {
      const uint32_t S0 = s * kSeg;
      const uint32_t S1 = min(S0 + kSeg, n);
      const uint32_t E = min(mlimit, S1);
      const uint32_t pend = min(S1, npos);
      const uint32_t cnt = pend > S0 ? pend - S0 : 0u;

      // ---- 1. a record per position, 4 consecutive positions per lane (128 per step): positions whose 4
      //         bytes repeat 4 / 8 / 1 bytes earlier are settled here (record = 0x8000 | distance), the
      //         others keep their hash for the table (record = hash)
      {
        const uint32_t sh = (mis & 3u) * 8u;
        for (uint32_t w0 = 4u * lane; w0 < cnt; w0 += 128u) {
          const uint32_t p0 = S0 + w0;
          const uint32_t* wp = s_chunk32 + ((mis + p0) >> 2);
          const uint32_t wm2 = wp[-2], wm1 = wp[-1], wc0 = wp[0], wc1 = wp[1], wc2 = wp[2];
          const uint32_t ym2 = __funnelshift_r(wm2, wm1, sh), ym1 = __funnelshift_r(wm1, wc0, sh),
                         y0 = __funnelshift_r(wc0, wc1, sh), y1 = __funnelshift_r(wc1, wc2, sh);
          const bool v4 = p0 >= 4u, v8 = p0 >= 8u;
          uint32_t rec[4];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const uint32_t key = __funnelshift_r(y0, y1, 8 * u);
            const uint32_t k4 = __funnelshift_r(ym1, y0, 8 * u), k8 = __funnelshift_r(ym2, ym1, 8 * u);
            const uint32_t k1 = u == 0 ? __funnelshift_r(ym1, y0, 24) : __funnelshift_r(y0, y1, 8 * u - 8);
            uint32_t r = (key * 2654435761u) >> (32 - HL);
            if ((u > 0 || v4) && key == k1) r = 0x8001u; // p >= 1
            if (v8 && key == k8) r = 0x8008u;
            if (v4 && key == k4) r = 0x8004u;
            rec[u] = r;
          }
          *reinterpret_cast<uint2*>(db + w0) = make_uint2(rec[0] | (rec[1] << 16), rec[2] | (rec[3] << 16));
        }
      }
      __syncwarp();

      // ---- 2. table probe + insert, in segment order across the team (the turn chain).  A window of 32
      //         positions reads the table before any of its lanes writes; of the lanes that share a hash the
      //         highest position stays in the table (stores repeated until it does).
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
          const bool tbl = !(rec[u] & 0x8000u);
          const uint32_t pp = S0 + w0 + 32 * u + lane;
          uint16_t* slot = s_table + (rec[u] & 0x7FFFu);
          uint32_t t = 0xFFFFu;
          if (tbl) {
            t = *slot;
            *slot = (uint16_t)pp;
          }
          __syncwarp();
          bool lost = tbl && (uint32_t)*reinterpret_cast<volatile uint16_t*>(slot) < pp;
          while (__any_sync(kFull, lost)) {
            if (lost) *slot = (uint16_t)pp;
            __syncwarp();
            lost = lost && (uint32_t)*reinterpret_cast<volatile uint16_t*>(slot) < pp;
          }
          rec[u] = tbl ? (t < pp ? pp - t : 0u) : (rec[u] & 0x7FFFu);
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

      // ---- 3. select: greedy walk over windows of 32 positions starting at the cursor.  The chosen match
      //         starts are collected as a bitmap (lane w: positions S0 + 32 w ..), their ends are left in the
      //         d[] entry behind the start (dead: it lies inside the match).
      uint32_t bm = 0;
      uint32_t lit = S0; // end of the last chosen match
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
        // proves [q, run end + 3) equal.  The lane that starts a run compares the next 4 bytes (which
        // settles the common case: the position behind the run failed at the same distance); bit 31 marks
        // an end that is still open after that.  Every lane of the run then takes the run's end.
        const uint32_t dprev = __shfl_up_sync(kFull, d, 1);
        const bool cont = ok && lane > 0u && dprev == d && ((vmask >> (lane - 1u)) & 1u);
        const unsigned cmask = __ballot_sync(kFull, cont);
        uint32_t mend = 0;
        if (ok && !cont) {
          const unsigned above = ~cmask & (0xFFFFFFFEu << lane);
          uint32_t a = c + 3u + (above ? (uint32_t)__ffs((int)above) - 1u : 32u);
          const uint32_t x = team::lds_u32_unaligned(chunk_sa + a) ^ team::lds_u32_unaligned(chunk_sa + a - d);
          if (x) {
            mend = min(a + (((uint32_t)__ffs((int)x) - 1u) >> 3), E);
          } else {
            a += 4u;
            mend = a >= E ? E : (a | 0x80000000u);
          }
        }
        mend = __shfl_sync(kFull, mend, 31 - __clz((int)(~cmask & ~(0xFFFFFFFEu << lane))));
        unsigned m = vmask, chosen = 0;
        for (;;) {
          const int ql = __ffs((int)m) - 1;
          const uint32_t me = __shfl_sync(kFull, mend, ql);
          uint32_t mq = me & 0x7FFFFFFFu;
          if (me >> 31) { // long match: the warp compares 128 bytes per step
            const uint32_t dq = __shfl_sync(kFull, d, ql);
            for (uint32_t a0 = mq;; a0 += 128u) {
              const uint32_t a = a0 + 4u * lane;
              uint32_t x = 1u;
              if (a < E) x = team::lds_u32_unaligned(chunk_sa + a) ^ team::lds_u32_unaligned(chunk_sa + a - dq);
              const unsigned diff = __ballot_sync(kFull, x != 0u);
              if (diff) {
                const int fl = __ffs((int)diff) - 1;
                const uint32_t xx = __shfl_sync(kFull, x, fl);
                mq = min(a0 + 4u * (uint32_t)fl + (((uint32_t)__ffs((int)xx) - 1u) >> 3), E);
                break;
              }
            }
            if ((int)lane == ql) mend = mq;
          }
          chosen |= 1u << ql;
          lit = mq;
          const uint32_t rel = mq - c;
          if (rel >= 32u) break;
          m = vmask & (0xFFFFFFFFu << rel);
          if (!m) break;
        }
        if ((chosen >> lane) & 1u) team::sts_u16(db_sa + 2u * (p - S0) + 2u, mend - S0);
        {
          const uint32_t rel0 = c - S0, wi = rel0 >> 5, shl = rel0 & 31u;
          if (lane == wi) bm |= chosen << shl;
          if (lane == wi + 1u) bm |= __funnelshift_l(chosen, 0u, shl);
        }
        c = max(lit, c + 32u);
      }
      __syncwarp();

      // ---- 4. sizes.  Lane w owns the sequences that start in its 32 positions, in order; the first
      //         sequence of the segment is sized and written apart (its literals begin in an earlier
      //         segment, or it merges with the pending match).
      const unsigned ne = __ballot_sync(kFull, bm != 0u);
      const bool have_first = ne != 0u;
      uint32_t f_q = 0, f_d = 0, f_mend = 0, l_d = 0, l_mend = 0, l_q = 0;
      uint32_t prev_end = S0; // end of the sequence in front of this lane's first one
      if (have_first) {
        uint32_t first_q = 0, first_d = 0, first_end = 0, last_q = 0, last_d = 0, last_end = 0;
        if (bm) {
          const uint32_t lo = 32u * lane + (uint32_t)__ffs((int)bm) - 1u, hi = 32u * lane + 31u - (uint32_t)__clz((int)bm);
          first_q = S0 + lo, first_d = db[lo], first_end = S0 + db[lo + 1u];
          last_q = S0 + hi, last_d = db[hi], last_end = S0 + db[hi + 1u];
        }
        const int fl = __ffs((int)ne) - 1, ll = 31 - __clz((int)ne);
        f_q = __shfl_sync(kFull, first_q, fl), f_d = __shfl_sync(kFull, first_d, fl), f_mend = __shfl_sync(kFull, first_end, fl);
        l_q = __shfl_sync(kFull, last_q, ll), l_d = __shfl_sync(kFull, last_d, ll), l_mend = __shfl_sync(kFull, last_end, ll);
        const unsigned lower = ne & lt_mask;
        const uint32_t pe = __shfl_sync(kFull, last_end, lower ? 31 - __clz((int)lower) : 0);
        if (lower) prev_end = pe;
      }
      const uint32_t f_mlen = f_mend - f_q;
      const bool l_open = have_first && l_mend == S1; // only the last match can stop at S1
      const bool l_first = l_q == f_q;
      const bool owns_first = have_first && (int)lane == __ffs((int)ne) - 1;
      uint32_t lane_bytes = 0, last_size = 0, last_hi = 0;
      {
        uint32_t b = bm, pe = prev_end;
        bool skip = owns_first;
        while (b) {
          const uint32_t i = 32u * lane + (uint32_t)__ffs((int)b) - 1u;
          b &= b - 1u;
          const uint32_t q = S0 + i, me = S0 + db[i + 1u];
          uint32_t size = 0;
          if (!skip) {
            const uint32_t nlit = q - pe, ml = me - q - 4u;
            size = 3u + nlit + (nlit >= 15u ? (nlit - 15u) / 255u + 1u : 0u) + ((me != S1 && ml >= 15u) ? (ml - 15u) / 255u + 1u : 0u);
            last_hi = nlit >= 15u ? 15u : nlit;
          }
          skip = false;
          lane_bytes += size;
          last_size = size;
          pe = me;
        }
      }
      uint32_t rest_bytes; // everything behind the first sequence
      const uint32_t lane_off = warp_excl_sum(lane_bytes, rest_bytes);
      // token of the last sequence, relative to the end of the first one
      const uint32_t l_tok_rel = __shfl_sync(kFull, lane_off + lane_bytes - last_size, have_first ? 31 - __clz((int)ne) : 0);
      const uint32_t l_hi = __shfl_sync(kFull, last_hi, have_first ? 31 - __clz((int)ne) : 0);

      // ---- 5. behind the predecessor's cursor: settle the pending match and the first sequence, pass the
      //         cursor on, then write the rest
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
      TeamState ns;       // state handed to the next segment
      bool whole = false; // the segment is one continuation of the pending match, which stays open
      bool merged = false; // the first sequence is part of the pending match
      bool close_pending = false;
      uint32_t pend_total = 0;
      uint32_t lit_from = st.value;
      if (st.flags & 1u) {
        const uint32_t o_d = st.flags >> 16;
        pend_total = st.value;
        merged = have_first && f_q == S0;
        if (merged && f_d != o_d) {
          // the pending match's own offset must reproduce this whole match
          for (uint32_t k = 0; k < f_mlen && merged; k += 128u) {
            const uint32_t a = S0 + k + 4u * lane;
            uint32_t x = 0;
            if (a < S0 + f_mlen) {
              x = team::lds_u32_at(s_chunk32, mis + a) ^ team::lds_u32_at(s_chunk32, mis + a - o_d);
              const uint32_t left = S0 + f_mlen - a;
              if (left < 4u) x &= (1u << (8u * left)) - 1u;
            }
            if (__any_sync(kFull, x != 0u)) merged = false;
          }
        }
        if (merged) pend_total += f_mlen;
        if (merged && l_open && l_first) {
          whole = true;
        } else {
          close_pending = true;
          lit_from = S0;
        }
      }
      const uint32_t pend_lsic_at = ob;
      if (close_pending && pend_total - 4u >= 15u) ob += (pend_total - 19u) / 255u + 1u;
      const uint32_t first_at = ob; // where the first sequence's token goes (when it is written)
      uint32_t first_nlit = 0;
      const bool write_first = have_first && !merged;
      const bool first_open = l_open && l_first;
      if (write_first) {
        first_nlit = f_q - lit_from;
        ob += 3u + first_nlit + (first_nlit >= 15u ? (first_nlit - 15u) / 255u + 1u : 0u) +
              ((!first_open && f_mlen - 4u >= 15u) ? (f_mlen - 19u) / 255u + 1u : 0u);
      }
      const uint32_t rest_at = ob;
      if (whole) {
        ns.out_base = ob;
        ns.flags = st.flags;
        ns.value = pend_total;
        ns.tok = st.tok;
      } else if (!have_first) {
        ns.out_base = ob;
        ns.flags = 0;
        ns.value = lit_from;
        ns.tok = 0;
      } else {
        ns.out_base = rest_at + rest_bytes;
        if (l_open) {
          ns.flags = 1u | (l_d << 16);
          ns.value = l_mend - l_q;
          const uint32_t fhi = first_nlit >= 15u ? 15u : first_nlit;
          ns.tok = l_first ? (first_at | (fhi << 20)) : ((rest_at + l_tok_rel) | (l_hi << 20));
        } else {
          ns.flags = 0;
          ns.value = lit;
          ns.tok = 0;
        }
      }
      if (s + 1 < nseg) {
        if (lane == 0) {
          s_state[nextw] = make_uint4(ns.out_base, ns.flags, ns.value, ns.tok);
          team::mbar_arrive(pub_bar + 2 * nextw);
        }
      }

      // ---- 6. write (off the hand-over chain)
      if (close_pending) {
        const uint32_t m = pend_total - 4u;
        if (lane == 0) gout[st.tok & 0xFFFFFu] = (uint8_t)(((st.tok >> 20) << 4) | (m >= 15u ? 15u : m));
        if (m >= 15u) team::put_lsic(gout + pend_lsic_at, m - 15u);
      }
      if (write_first) {
        uint32_t o = first_at;
        const uint32_t fm = f_mlen - 4u;
        // the token of a sequence whose match is still open is written by whoever closes it
        if (lane == 0 && !first_open) gout[o] = (uint8_t)(((first_nlit >= 15u ? 15u : first_nlit) << 4) | (fm >= 15u ? 15u : fm));
        o += 1u;
        if (first_nlit >= 15u) o += team::put_lsic(gout + o, first_nlit - 15u);
        team::copy_s2g(gout + o, s_chunk32, mis + lit_from, first_nlit);
        o += first_nlit;
        if (lane == 0) {
          gout[o] = (uint8_t)(f_d & 0xffu);
          gout[o + 1] = (uint8_t)(f_d >> 8);
        }
        o += 2u;
        if (!first_open && fm >= 15u) team::put_lsic(gout + o, fm - 15u);
      }
      {
        uint32_t b = bm, pe = prev_end;
        bool skip = owns_first;
        uint8_t* o = gout + rest_at + lane_off;
        uint32_t long_src = 0, long_cnt = 0;
        uint8_t* long_dst = nullptr;
        while (b) {
          const uint32_t i = 32u * lane + (uint32_t)__ffs((int)b) - 1u;
          b &= b - 1u;
          const uint32_t q = S0 + i, dd = db[i], me = S0 + db[i + 1u];
          if (!skip) {
            const uint32_t nlit = q - pe, ml = me - q - 4u;
            const bool open = me == S1;
            if (!open) *o = (uint8_t)(((nlit >= 15u ? 15u : nlit) << 4) | (ml >= 15u ? 15u : ml));
            ++o;
            if (nlit >= 15u) {
              uint32_t x = nlit - 15u;
              for (; x >= 255u; x -= 255u) *o++ = 0xff;
              *o++ = (uint8_t)x;
            }
            if (nlit <= 28u) {
              for (uint32_t k = 0; k < nlit; ++k) o[k] = s_chunk[mis + pe + k];
            } else { // at most one per lane: later sequences of the lane have fewer than 28 literals
              long_src = pe, long_cnt = nlit, long_dst = o;
            }
            o += nlit;
            o[0] = (uint8_t)(dd & 0xffu);
            o[1] = (uint8_t)(dd >> 8);
            o += 2;
            if (!open && ml >= 15u) {
              uint32_t x = ml - 15u;
              for (; x >= 255u; x -= 255u) *o++ = 0xff;
              *o++ = (uint8_t)x;
            }
          }
          skip = false;
          pe = me;
        }
        // long literal runs: the whole warp copies them, one run at a time
        unsigned longm = __ballot_sync(kFull, long_cnt != 0u);
        while (longm) {
          const int l = __ffs((int)longm) - 1;
          longm &= longm - 1u;
          const uint32_t src = __shfl_sync(kFull, long_src, l), cntl = __shfl_sync(kFull, long_cnt, l);
          uint8_t* dst = reinterpret_cast<uint8_t*>(__shfl_sync(kFull, reinterpret_cast<unsigned long long>(long_dst), l));
          team::copy_s2g(dst, s_chunk32, mis + src, cntl);
        }
      }
      if (s + 1 == nseg) {
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
