// b200comp — LZ4 block encoder for chunks <= 64 KiB: one CTA ("team" of W warps) per chunk.
//
// Replaces the reference's one-warp-per-chunk greedy encoder (src/LZ4Kernels.hiph:793-969,
// launched by src/lowlevel/LZ4CompressionKernels.hip:73-94), whose per-window chain
// "key -> table probe -> verify -> extend" is serial per chunk.  The output is a plain LZ4 block
// (decodable by the reference decoder and by liblz4) but not the reference's byte stream; the
// policy is stated sequentially in oracle/lz_team_model.c, which the tests compare against byte
// for byte.
//
// Data movement: the chunk is staged in shared memory by one bulk copy (cp.async.bulk +
// mbarrier), the hash table (32-bit entries, position + 1) lives in shared memory, neither touches
// HBM again.  The chunk is cut into segments of 1024 positions, dealt round-robin to the warps.
// Per segment a warp
//   1. hashes its positions (parallel across warps; 4/8/1-byte repeats are settled here),
//   2. takes the "turn" (an mbarrier chain in segment order): per window of 32 positions one table
//      load and one shared-memory atomic max, nothing else is serialised per chunk.  It yields
//      d[p] = distance to the most recent earlier position with the same hash,
//   3. verifies every candidate (4 bytes at p and p - d[p]); window u of the segment is lane u's block
//      and the window's ballot is that lane's candidate bitmap.  Then every LANE parses its block of
//      32 positions greedily on its own from the block's first position (next candidate, extend
//      4 bytes per step); the matches of a lane (at most 8) go to a small per-lane list,
//   4. X chain: waits for X = the largest match end in front of the segment, settles "long" matches
//      with the whole warp, passes X on at once; off the chain it cuts every lane's matches against X
//      (an exclusive prefix maximum over the lanes) and sizes the sequences; output chain: waits for
//      (output cursor, first unwritten literal), adds the one literal run that depends on it, passes on,
//   5. writes its sequences to global memory, every lane its own, long literal runs by the warp.
#pragma once

#include "lz_team.cuh"

namespace b200comp {
namespace team {

constexpr uint32_t kLongMatch = 1024; // a match longer than this is written by the whole warp
constexpr uint32_t kLaneSteps = 15; // further 4-byte words a lane verifies behind the first 4 bytes of its match

template <int CAP, int HL, int W>
struct TeamLayout {
  static constexpr uint32_t kFront = 16;          // readable bytes in front of the chunk (phase 1 looks 8 back)
  static constexpr uint32_t kChunkBuf = kFront + CAP + 48; // + 16 bytes of phase slack + 32 bytes of over-read pad
  static constexpr uint32_t kTableOff = kChunkBuf;
  static constexpr uint32_t kDbufOff = kTableOff + (4u << HL);
  static constexpr uint32_t kEndsOff = kDbufOff + W * kSeg * 2;   // per lane 8 match ends (uint16)
  static constexpr uint32_t kCtrlOff = kEndsOff + W * 32 * 16;
  static constexpr uint32_t kDummyOff = kCtrlOff + 16 + 80 * W; // load barrier; per warp: turn, X and output barriers + their two states
  static constexpr uint32_t kBytes = kDummyOff + 128 * W;     // a word per lane for the lanes without a table slot
};

template <class Emit, int CAP, int HL, int W>
__global__ void __launch_bounds__(W * 32, (CAP > 16384 ? 1 : (CAP > 4096 ? 3 : 6)))
compress_team_kernel(const void* const* in_ptrs, const size_t* in_bytes, void* const* out_ptrs, size_t* out_bytes,
                     size_t batch)
{
  using L = TeamLayout<CAP, HL, W>;
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* s_chunk = smem + L::kFront;
  const uint32_t* s_chunk32 = reinterpret_cast<const uint32_t*>(smem + L::kFront);
  uint32_t* s_table = reinterpret_cast<uint32_t*>(smem + L::kTableOff);
  uint16_t* s_dbuf = reinterpret_cast<uint16_t*>(smem + L::kDbufOff);
  uint16_t* s_ends = reinterpret_cast<uint16_t*>(smem + L::kEndsOff);
  uint64_t* load_bar = reinterpret_cast<uint64_t*>(smem + L::kCtrlOff);
  uint64_t* turn_bar = load_bar + 2;          // [W], one 16-byte slot each
  uint64_t* pub_bar = turn_bar + 2 * W;       // [W]
  uint64_t* x_bar = pub_bar + 2 * W;          // [W]
  uint4* s_state = reinterpret_cast<uint4*>(x_bar + 2 * W); // [W] output cursor + first unwritten literal
  uint4* s_xstate = s_state + W;              // [W] X = largest match end so far

  const uint32_t lane = (uint32_t)lane_id();
  const uint32_t warp = threadIdx.x >> 5;
  const unsigned lt_mask = lanemask_lt();

  if (threadIdx.x == 0) {
    mbar_init(load_bar, 1);
    for (int w = 0; w < W; ++w) {
      mbar_init(turn_bar + 2 * w, 1);
      mbar_init(pub_bar + 2 * w, 1);
      mbar_init(x_bar + 2 * w, 1);
    }
    mbar_init_fence();
  }
  uint32_t load_phase = 0, turn_phase = 0, pub_phase = 0, x_phase = 0;

  for (size_t chunk = blockIdx.x; chunk < batch; chunk += gridDim.x) {
    __syncthreads(); // the previous chunk is finished with shared memory (and the barriers are initialised)
    const uint8_t* src = static_cast<const uint8_t*>(in_ptrs[chunk]);
    uint8_t* gout = static_cast<uint8_t*>(out_ptrs[chunk]);
    const size_t n64 = in_bytes[chunk];
    if (n64 == 0 || n64 > (size_t)CAP) { // empty chunk -> the codec's empty stream; larger than the caller's declared maximum -> 0
      if (threadIdx.x == 0) out_bytes[chunk] = n64 == 0 ? Emit::empty_stream(gout) : 0;
      continue;
    }
    const uint32_t n = (uint32_t)n64;
    const uint32_t mis = stage_chunk(s_chunk, src, n, load_bar);
    {
      uint4* t128 = reinterpret_cast<uint4*>(s_table);
      const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
      for (uint32_t i = threadIdx.x; i < (4u << HL) / 16u; i += W * 32) t128[i] = zero;
    }
    mbar_wait(load_bar, load_phase);
    load_phase ^= 1u;
    __syncthreads();

    const uint32_t npos = n >= 12u ? n - 11u : 0u;  // positions that may start a match: p <= n - 12
    const uint32_t E = n >= 5u ? n - 5u : 0u;       // matches end at or before n - 5
    const uint32_t nseg = (n + kSeg - 1u) / kSeg;
    uint16_t* db = s_dbuf + warp * kSeg;
    uint16_t* ends = s_ends + warp * 256u;          // [k * 32 + lane]
    const uint32_t nextw = (warp + 1u) % W;
    const uint32_t chunk_sa = smem_addr(s_chunk) + mis; // shared address of chunk byte 0
    const uint32_t db_sa = smem_addr(db);
    const uint32_t ends_sa = smem_addr(ends) + 2u * lane;

    for (uint32_t s = warp; s < nseg; s += W) {
      const uint32_t S0 = s * kSeg;
      const uint32_t S1 = min(S0 + kSeg, n);
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
      //         positions reads the table before any of its lanes inserts; the insert is an atomic max, so the
      //         table holds the highest position per hash whatever the order inside the window.
      uint32_t cand = 0; // positions of the lane's block whose candidate holds
      {
        const uint32_t table_sa = smem_addr(s_table);
        const uint32_t dummy_sa = smem_addr(smem + L::kDummyOff) + 128u * warp + 4u * lane;
        uint32_t sa[32]; // shared address of the lane's table slot per window (or its dummy word); then the slot's old value
#pragma unroll
        for (int u = 0; u < 32; ++u) {
          const uint32_t i = 32u * u + lane;
          const uint32_t rec = i < cnt ? (uint32_t)db[i] : 0x8000u;
          sa[u] = (rec & 0x8000u) ? dummy_sa : table_sa + 4u * rec;
        }
        if (s > 0) {
          mbar_wait(turn_bar + 2 * warp, turn_phase);
          turn_phase ^= 1u;
        }
        __syncwarp();
        // The serial section of the chunk: two shared-memory instructions per window.  The warp is
        // converged here and its shared-memory instructions execute in program order, which is what orders the
        // loads of a window before its inserts and those before the next window's loads.
        const uint32_t pp1 = S0 + lane + 1u;
#pragma unroll
        for (int u = 0; u < 32; ++u) {
          const uint32_t a = sa[u];
          sa[u] = lds_u32_ordered(a);
          red_max_u32(a, pp1 + 32u * u);
          if ((u & 7) == 7) __syncwarp();
        }
        if (s + 1 < nseg) {
          if (lane == 0) mbar_arrive(turn_bar + 2 * nextw);
        }
        // distances, and which candidates hold: the 4 bytes at p and at p - d[p] are equal.  Window u is the
        // block of lane u, which keeps the window's ballot as its candidate bitmap.
#pragma unroll
        for (int u = 0; u < 32; ++u) {
          const uint32_t i = 32u * u + lane;
          bool ok = false;
          if (i < cnt) {
            const uint32_t rec = db[i];
            const uint32_t t = sa[u];
            const uint32_t d = (rec & 0x8000u) ? (rec & 0x7FFFu) : (t ? pp1 + 32u * u - t : 0u);
            db[i] = (uint16_t)d;
            const uint32_t pa = chunk_sa + S0 + i;
            if (d) ok = lds_u32_unaligned(pa) == lds_u32_unaligned(pa - d);
          }
          const unsigned wm = __ballot_sync(kFull, ok);
          if ((int)lane == u) cand = wm;
        }
      }
      __syncwarp();

      // ---- 3. parse: lane l owns positions S0 + 32 l .. S0 + 32 l + 31 and walks them on its own: the next
      //         candidate at or behind the cursor starts a match, which is extended one 4-byte comparison per
      //         step; the cursor moves to its end.
      //         bm = starts of its matches (bit = position in the block), ends in the lane's list,
      //         is_long = the last match was still equal after kLaneSteps words (its end is open).
      uint32_t bm = 0, nm = 0, v = 0;
      bool is_long = false;
      const uint32_t B0 = S0 + 32u * lane;
      {
        uint32_t q = 0, m = 0, d = 0, steps = 0;
        bool ext = false;
        while (ext || cand) {
          if (!ext) {
            const uint32_t bit = (uint32_t)__ffs((int)cand) - 1u;
            q = B0 + bit;
            d = lds_u16(db_sa + 2u * (q - S0));
            m = q + 4u, steps = 0, ext = true; // m < E: q <= n - 12
          }
          const uint32_t x = lds_u32_unaligned(chunk_sa + m) ^ lds_u32_unaligned(chunk_sa + m - d);
          bool done = true;
          uint32_t end;
          if (x) {
            end = min(m + (((uint32_t)__ffs((int)x) - 1u) >> 3), E);
          } else {
            m += 4u;
            ++steps;
            end = m;
            if (m >= E) {
              end = E;
            } else if (steps == kLaneSteps) {
              is_long = true;
            } else {
              done = false;
            }
          }
          if (done) {
            bm |= 1u << (q - B0);
            sts_u16(ends_sa + 64u * nm, end);
            ++nm;
            v = end;
            ext = false;
            const uint32_t rel = end - B0;
            cand = rel >= 32u ? 0u : cand & (0xFFFFFFFFu << rel);
          }
        }
      }
      __syncwarp();

      // ---- 4a. X chain: X = the largest match end in front of the segment.  Long matches are settled here, in
      //          lane order: dropped when the verified part ends at or before the X in front of them, else the
      //          warp compares 512 bytes per step up to the true end.
      uint32_t Xin = 0;
      if (s > 0) {
        mbar_wait(x_bar + 2 * warp, x_phase);
        x_phase ^= 1u;
        Xin = s_xstate[warp].x;
      }
      {
        // (a verified part that ends at or before the X in front of the segment is dropped without a look)
        unsigned longm = __ballot_sync(kFull, is_long && v > Xin);
        while (longm) {
          const int ll = __ffs((int)longm) - 1;
          longm &= longm - 1u;
          const uint32_t xl = max(Xin, __reduce_max_sync(kFull, (int)lane < ll ? v : 0u));
          const uint32_t vl = __shfl_sync(kFull, v, ll);
          if (vl > xl) {
            uint32_t dq = 0;
            if ((int)lane == ll) dq = db[31u - (uint32_t)__clz((int)bm) + 32u * lane];
            dq = __shfl_sync(kFull, dq, ll);
            uint32_t mq = E;
            for (uint32_t a0 = vl; a0 < E; a0 += 512u) { // 16 bytes per lane and step
              const uint32_t a = a0 + 16u * lane;
              uint32_t fd = 16u; // first byte of the lane's 16 that differs (or lies at / behind E)
#pragma unroll
              for (int k = 3; k >= 0; --k) {
                const uint32_t ak = a + 4u * (uint32_t)k;
                uint32_t x = 1u;
                if (ak < E) x = lds_u32_unaligned(chunk_sa + ak) ^ lds_u32_unaligned(chunk_sa + ak - dq);
                if (x) fd = 4u * (uint32_t)k + (((uint32_t)__ffs((int)x) - 1u) >> 3);
              }
              const unsigned diff = __ballot_sync(kFull, fd < 16u);
              if (diff) {
                const int fl = __ffs((int)diff) - 1;
                mq = min(a0 + 16u * (uint32_t)fl + __shfl_sync(kFull, fd, fl), E);
                break;
              }
            }
            if ((int)lane == ll) {
              v = mq;
              sts_u16(ends_sa + 64u * (nm - 1u), mq);
            }
            longm &= __ballot_sync(kFull, v > mq); // everything this match covers is dropped
          }
        }
      }
      if (s + 1 < nseg) {
        const uint32_t Xout = max(Xin, __reduce_max_sync(kFull, v));
        if (lane == 0) {
          s_xstate[nextw] = make_uint4(Xout, 0u, 0u, 0u);
          mbar_arrive(x_bar + 2 * nextw);
        }
      }
      // X in front of every lane: exclusive prefix maximum of the lanes' last match ends
      uint32_t X;
      {
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t t = __shfl_up_sync(kFull, incl, o);
          if ((int)lane >= o) incl = max(incl, t);
        }
        X = __shfl_up_sync(kFull, incl, 1);
        X = lane == 0 ? Xin : max(X, Xin);
      }

      // ---- 4b. every lane cuts its matches against X: the ones that keep fewer than 4 bytes are dropped (they
      //          are the first ones of the lane), the first survivor starts at max(start, X), the others are
      //          unchanged.  Sizes of everything but the segment's first literal run (it starts at the
      //          predecessor's last emitted match end, which arrives with the output cursor).
      uint32_t em = 0, kfirst = 0, sfirst = 0, bytes = 0, last_end = 0;
      {
        uint32_t b = bm, k = 0;
        bool found = false;
        while (b) {
          const uint32_t bit = (uint32_t)__ffs((int)b) - 1u;
          const uint32_t q = B0 + bit, e = lds_u16(ends_sa + 64u * k);
          if (!found) {
            const uint32_t s0 = max(q, X);
            if (e >= s0 + 4u) {
              found = true;
              sfirst = s0, kfirst = k, em = b;
              bytes = Emit::match_bytes(e - s0, Emit::kNeedDist ? (uint32_t)db[32u * lane + bit] : 0u);
              last_end = e;
            }
          } else {
            const uint32_t nlit = q - last_end;
            bytes += Emit::lit_bytes(nlit) + Emit::match_bytes(e - q, Emit::kNeedDist ? (uint32_t)db[32u * lane + bit] : 0u);
            last_end = e;
          }
          b &= b - 1u;
          ++k;
        }
      }
      const unsigned emit_mask = __ballot_sync(kFull, em != 0u);
      const unsigned lower = emit_mask & lt_mask;
      uint32_t Lf = __shfl_sync(kFull, last_end, lower ? 31 - __clz((int)lower) : 0); // where the lane's first literals begin
      const bool seg_first = em != 0u && lower == 0u;
      if (em != 0u && lower != 0u) {
        bytes += Emit::lit_bytes(sfirst - Lf);
      }
      uint32_t total;
      uint32_t lane_off = warp_excl_sum(bytes, total);
      const uint32_t seg_last_end = __shfl_sync(kFull, last_end, emit_mask ? 31 - __clz((int)emit_mask) : 0);
      const uint32_t seg_sfirst = __shfl_sync(kFull, sfirst, emit_mask ? __ffs((int)emit_mask) - 1 : 0);

      // ---- 4c. output chain: (output cursor, first literal not yet written)
      uint32_t ob = 0, Lin = 0;
      if (s == 0) ob = Emit::preamble(gout, n, lane);
      if (s > 0) {
        mbar_wait(pub_bar + 2 * warp, pub_phase);
        pub_phase ^= 1u;
        const uint4 st = s_state[warp];
        ob = st.x, Lin = st.y;
      }
      uint32_t Lout = Lin;
      if (emit_mask) {
        const uint32_t extra = Emit::lit_bytes(seg_sfirst - Lin);
        total += extra;
        Lout = seg_last_end;
        if (seg_first) Lf = Lin;
        else lane_off += extra;
      }
      if (s + 1 < nseg) {
        if (lane == 0) {
          s_state[nextw] = make_uint4(ob + total, Lout, 0u, 0u);
          mbar_arrive(pub_bar + 2 * nextw);
        }
      }

      // ---- 5. write (off the hand-over chain): every lane its own sequences
      {
        uint32_t b = em, k = kfirst, pe = Lf;
        bool first = true;
        uint8_t* o = gout + ob + lane_off;
        uint32_t long_src = 0, long_cnt = 0, lm_len = 0, lm_d = 0;
        uint8_t *long_dst = nullptr, *lm_dst = nullptr;
        while (b) {
          const uint32_t bit = (uint32_t)__ffs((int)b) - 1u;
          b &= b - 1u;
          const uint32_t q = first ? sfirst : B0 + bit;
          const uint32_t e = lds_u16(ends_sa + 64u * k), dd = db[32u * lane + bit];
          ++k;
          first = false;
          const uint32_t nlit = q - pe;
          o = Emit::put_lit_header(o, nlit, e - q);
          if (nlit <= 28u) {
            for (uint32_t i = 0; i < nlit; ++i) o[i] = s_chunk[mis + pe + i];
          } else { // at most one per lane: later sequences of the lane have fewer than 28 literals
            long_src = pe, long_cnt = nlit, long_dst = o;
          }
          o += nlit;
          if (e - q > kLongMatch) { // necessarily the lane's last match: written by the whole warp below
            lm_dst = o, lm_len = e - q, lm_d = dd;
          } else {
            o = Emit::put_match(o, dd, e - q);
          }
          pe = e;
        }
        // long literal runs: the whole warp copies them, one run at a time
        unsigned longm = __ballot_sync(kFull, long_cnt != 0u);
        while (longm) {
          const int l = __ffs((int)longm) - 1;
          longm &= longm - 1u;
          const uint32_t srcp = __shfl_sync(kFull, long_src, l), cntl = __shfl_sync(kFull, long_cnt, l);
          uint8_t* dst = reinterpret_cast<uint8_t*>(__shfl_sync(kFull, reinterpret_cast<unsigned long long>(long_dst), l));
          copy_s2g(dst, s_chunk32, mis + srcp, cntl);
        }
        // long matches (hundreds of length bytes / copy elements): likewise
        longm = __ballot_sync(kFull, lm_len != 0u);
        while (longm) {
          const int l = __ffs((int)longm) - 1;
          longm &= longm - 1u;
          const uint32_t len = __shfl_sync(kFull, lm_len, l), d = __shfl_sync(kFull, lm_d, l);
          uint8_t* dst = reinterpret_cast<uint8_t*>(__shfl_sync(kFull, reinterpret_cast<unsigned long long>(lm_dst), l));
          Emit::put_match_warp(dst, d, len, lane);
        }
      }
      if (s + 1 == nseg) {
        // last segment: the final literals
        uint32_t o2 = ob + total;
        const uint32_t nlit = n - Lout;
        o2 += Emit::put_tail_header(gout + o2, nlit, lane);
        copy_s2g(gout + o2, s_chunk32, mis + Lout, nlit);
        o2 += nlit;
        if (lane == 0) out_bytes[chunk] = o2;
      }
      __syncwarp();
    }
  }
}

} // namespace team
} // namespace b200comp
