// Body of the softmax warpgroups of prefill_tcgen05_kernel (textually included inside the kernel).
//
// BOTH warpgroups work on EVERY S tile: warp w and warp w + 4 own the same 32 TMEM lanes (query rows), warpgroup 0
// takes columns 0..63 of the tile and warpgroup 1 columns 64..127.  A thread therefore runs 64 ex2 per tile instead of
// 128 -- the MUFU unit retires one warp instruction per 8 clocks, so a full row per thread took >= 1024 clocks for the
// exponentials alone and the S -> P latency (2200 clocks measured) exceeded the ~1500 clocks the other tile's two MMA
// groups cover; the tensor pipe waited for P.  Halved, the softmax of tile A hides under the MMAs of tile B and vice
// versa.  The two halves of a row agree on the row max through shared memory and a 64-thread named barrier.
{
  constexpr int GPT = kSoftmaxGroups / kTeams;   // warpgroups (column groups) per team
  constexpr int CG = TN / GPT;                   // S / O columns of a row this thread owns
  const int team = (warp >> 2) / GPT;            // kTeams == 2: the query tile (A or B) this warpgroup works on
  const int grp = (warp >> 2) % GPT;             // column group inside the tile
  const int row = tid & 127;
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const float c = p.scale_log2;
  const int pair_bar = 1 + (warp & 3) + 4 * team;   // named barrier of the GPT warps sharing these lanes and this tile
  float* const xch = reinterpret_cast<float*>(smem_gen + kSmemXchg) + team * (2 * GPT * 128);   // [2 parities][GPT][128 rows]
  uint32_t x_cnt = 0;
  uint32_t s_cnt[2] = {0, 0};  // S tiles consumed per query tile (phase of s_full[t])
  uint32_t o_cnt[2] = {0, 0};  // O commits consumed (phase of o_full[t]); one per PV, waited exactly once each

  // values of the partner threads (same row, other column groups): max or sum over all of them, own value included
  auto exchange = [&](float v, bool take_max) -> float {
    float* xb = xch + (x_cnt & 1u) * (GPT * 128);
    ++x_cnt;
    xb[grp * 128 + row] = v;
    asm volatile("bar.sync %0, %1;\n" ::"r"(pair_bar), "n"(32 * GPT) : "memory");
    float r = v;
#pragma unroll
    for (int g = 1; g < GPT; ++g) {
      const float o = xb[((grp + g) % GPT) * 128 + row];
      r = take_max ? fmaxf(r, o) : r + o;
    }
    return r;
  };

  int item_k = 0;
  // the next item's descriptor (four dependent-free global loads) is fetched while the current item's epilogue runs
  Item nxt = decode_item(p, blockIdx.x, num_pairs, batch);   // gridDim.x <= total_items
  for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
    const Item it = nxt;
    const int w_next = w + static_cast<int>(gridDim.x);
    if (!it.live) {
      if (w_next < total_items) nxt = decode_item(p, w_next, num_pairs, batch);
      continue;
    }
    const bool tri = kDebug && debug != nullptr && blockIdx.x == 0 && tid == 0;
    if (tri) item_stamp(debug, item_k, 0);
    float m_used[2] = {-INFINITY, -INFINITY};  // max the stored P/O/l are relative to (may lag the true max by < 2^8)
    float l_run[2] = {0.f, 0.f};               // this thread's columns only
    const bool tr = kDebug && debug != nullptr && blockIdx.x == 0 && w == static_cast<int>(blockIdx.x) && tid == 0;

    for (int j = 0; j < it.n_tiles; ++j) {
      const int t_base = j * TN + grp * CG;
#pragma unroll
      for (int tt = 0; tt < 2 / kTeams; ++tt) {
        const int t = kTeams == 2 ? team : tt;   // state arrays are indexed by tt (compile-time), tiles by t
        if (j >= it.n_t[t]) continue;
        const uint32_t s_addr = tmem + lane_base + static_cast<uint32_t>(t * TN + grp * CG);
        const uint32_t p_addr = tmem + lane_base + static_cast<uint32_t>(t * TN + grp * (CG / 2));  // bf16 pairs over S_t
        const uint32_t o_addr = tmem + lane_base + 2u * TN + static_cast<uint32_t>(t * kHeadDim + grp * CG);
        const int qrow = it.mr[t] + row;              // row inside the sequence
        const int limit = min(it.shift + qrow, it.Lk - 1);  // last visible key of this row
        const bool dump = kDebug && debug != nullptr && it.m0 == 0 && it.h == 0 && it.seq == 0 && t == 0;
        if (tr) trace_stamp(debug, 1 + t, j, 0);
        tc::mbar_wait(bar(kSFull + t), s_cnt[tt] & 1u);
        ++s_cnt[tt];
        tc::fence_after_sync();
        if (tr) trace_stamp(debug, 1 + t, j, 1);
        if (tri && j == 0 && tt == 0) item_stamp(debug, item_k, 1);
        uint32_t s[CG];
        if (dbg_mode & 2u) {
#pragma unroll
          for (int i = 0; i < CG; ++i) s[i] = 0u;
        } else {
#pragma unroll
          for (int cc = 0; cc < CG / 32; ++cc) tc::tmem_ld32p(s_addr + cc * 32, s + cc * 32);
          tc::tmem_ld_wait();
        }
        if (tr) trace_stamp(debug, 1 + t, j, 2);
        if (dump && j == 0) {
#pragma unroll
          for (int i = 0; i < CG; ++i) debug[row * TN + grp * CG + i] = __uint_as_float(s[i]);
        }
        const bool need_mask = (j * TN + TN - 1 > it.shift + it.mr[t]) || (j * TN + TN > it.Lk);
        if (need_mask) {
#pragma unroll
          for (int i = 0; i < CG; ++i)
            if (t_base + i > limit) s[i] = 0xff800000u;  // -inf
        }
        // four independent chains of three-input max (FMNMX3)
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
        for (int i = 0; i < CG; i += 8) {
          mx0 = fmax3(mx0, __uint_as_float(s[i]), __uint_as_float(s[i + 1]));
          mx1 = fmax3(mx1, __uint_as_float(s[i + 2]), __uint_as_float(s[i + 3]));
          mx2 = fmax3(mx2, __uint_as_float(s[i + 4]), __uint_as_float(s[i + 5]));
          mx3 = fmax3(mx3, __uint_as_float(s[i + 6]), __uint_as_float(s[i + 7]));
        }
        const float mx_half = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
        // The barrier inside also orders the partner's S loads (done: its values are in registers) before this
        // thread's P stores below, which overwrite S columns the partner owns (P aliases the first half of S_t).
        const float mx = exchange(mx_half, true);

        if (j == 0) {
          m_used[tt] = mx;
        } else {
          // PV(j-1) finished before S(j) could be produced (in-order tensor pipe): this wait only keeps phases in step
          tc::mbar_wait(bar(kOFull + t), o_cnt[tt] & 1u);
          ++o_cnt[tt];
          tc::fence_after_sync();
          const bool need = mx * c > m_used[tt] * c + kRescaleThreshold;   // same verdict in both halves of the row
          if (__any_sync(0xffffffffu, need)) {
            const float alpha = need ? fast_exp2(m_used[tt] * c - mx * c) : 1.f;
            if (need) m_used[tt] = mx;
            l_run[tt] *= alpha;
#pragma unroll
            for (int cc = 0; cc < CG / 32; ++cc) {
              uint32_t r[32];
              tc::tmem_ld32(o_addr + cc * 32, r);
              tc::tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
              tc::tmem_st32(o_addr + cc * 32, r);
            }
          }
        }
        if (tr) trace_stamp(debug, 1 + t, j, 3);
        const float m_sc = (m_used[tt] == -INFINITY) ? 0.f : m_used[tt] * c;
        const float2 c2 = make_float2(c, c);
        const float2 nm2 = make_float2(-m_sc, -m_sc);
        float2 psum = make_float2(0.f, 0.f);
        if (!(dbg_mode & 1u)) {
          if (need_mask || kEmu == 0) {   // masked tiles (one or two per item) hold -inf: MUFU only
#pragma unroll
            for (int cc = 0; cc < CG / 32; ++cc) {
              uint32_t pk[16];
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const float2 x = __ffma2_rn(make_float2(__uint_as_float(s[cc * 32 + 2 * i]),
                                                        __uint_as_float(s[cc * 32 + 2 * i + 1])), c2, nm2);
                const float2 pv = make_float2(fast_exp2(x.x), fast_exp2(x.y));
                psum = __fadd2_rn(psum, pv);
                pk[i] = pack_bf16x2(pv.x, pv.y);
              }
              tc::tmem_st16(p_addr + cc * 16, pk);
            }
          } else {
#pragma unroll
            for (int cc = 0; cc < CG / 32; ++cc) {
              uint32_t pk[16];
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const float2 x = __ffma2_rn(make_float2(__uint_as_float(s[cc * 32 + 2 * i]),
                                                        __uint_as_float(s[cc * 32 + 2 * i + 1])), c2, nm2);
                // pairs on the polynomial are spread between the MUFU ones so both pipes stay busy
                const bool poly = ((i & 7) * kEmu) / 8 != (((i & 7) + 1) * kEmu) / 8;
                const float2 pv = poly ? exp2_poly2(x) : make_float2(fast_exp2(x.x), fast_exp2(x.y));
                psum = __fadd2_rn(psum, pv);
                pk[i] = pack_bf16x2(pv.x, pv.y);
              }
              tc::tmem_st16(p_addr + cc * 16, pk);
            }
          }
        }
        if (tr) trace_stamp(debug, 1 + t, j, 4);
        tc::tmem_st_wait();
        tc::fence_before_sync();
        tc::mbar_arrive(bar(kPFull + t));
        if (tr) trace_stamp(debug, 1 + t, j, 5);
        l_run[tt] += psum.x + psum.y;
      }
    }

    if (tri) item_stamp(debug, item_k, 2);
    if (w_next < total_items) nxt = decode_item(p, w_next, num_pairs, batch);
    // ---- epilogue: O_t (TMEM) / l -> bf16 -> global; rows beyond the sequence are not stored
#pragma unroll
    for (int tt = 0; tt < 2 / kTeams; ++tt) {
      const int t = kTeams == 2 ? team : tt;
      const int nt = it.n_t[t];
      if (nt == 0) continue;
      const uint32_t o_addr = tmem + lane_base + 2u * TN + static_cast<uint32_t>(t * kHeadDim + grp * CG);
      const int qrow = it.mr[t] + row;
      const bool dump = kDebug && debug != nullptr && it.m0 == 0 && it.h == 0 && it.seq == 0 && t == 0;
      const float l_all = exchange(l_run[tt], false);
      tc::mbar_wait(bar(kOFull + t), o_cnt[tt] & 1u);
      ++o_cnt[tt];
      tc::fence_after_sync();
      const float inv = l_all > 0.f ? 1.f / l_all : 0.f;
      uint32_t o[CG];
#pragma unroll
      for (int cc = 0; cc < CG / 32; ++cc) tc::tmem_ld32p(o_addr + cc * 32, o + cc * 32);
      tc::tmem_ld_wait();
      tc::fence_before_sync();
      tc::mbar_arrive(bar(kOEmpty + t));  // the next item's first PV may overwrite O_t now
      if (tri && tt == 2 / kTeams - 1) item_stamp(debug, item_k, 3);
      if (dump && nt == 1) {
#pragma unroll
        for (int i = 0; i < CG; ++i) debug[TM * TN + row * kHeadDim + grp * CG + i] = __uint_as_float(o[i]);
      }
      // Output rows are 256 B apart per head: a thread storing its own row would make every store instruction touch 32
      // lines (measured: ~2000 clocks per tile, the largest part of the gap between items).  Each warp transposes
      // 32 rows x 32 columns at a time through its private staging rows (80 B pitch: conflict-free 16-byte writes), so a
      // store instruction covers 8 rows x 64 contiguous bytes.
      unsigned char* const stg = smem_gen + kSmemStage + warp * (32 * kStagePitch);
      const int r0 = it.mr[t] + (warp & 3) * 32;      // first row of this warp inside the sequence
#pragma unroll
      for (int pass = 0; pass < CG / 32; ++pass) {
#pragma unroll
        for (int c4 = 0; c4 < 4; ++c4) {
          const int i = pass * 32 + c4 * 8;
          uint4 v;
          v.x = pack_bf16x2(__uint_as_float(o[i + 0]) * inv, __uint_as_float(o[i + 1]) * inv);
          v.y = pack_bf16x2(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
          v.z = pack_bf16x2(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
          v.w = pack_bf16x2(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
          *reinterpret_cast<uint4*>(stg + lane * kStagePitch + c4 * 16) = v;
        }
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int r = i * 8 + (lane >> 2);
          const uint4 v = *reinterpret_cast<const uint4*>(stg + r * kStagePitch + (lane & 3) * 16);
          if (r0 + r < it.Lq) {
            __nv_bfloat16* dst = p.out + (static_cast<int64_t>(it.q0 + r0 + r) * p.num_q_heads + it.hq[t]) * kHeadDim +
                                 grp * CG + pass * 32 + (lane & 3) * 8;
            *reinterpret_cast<uint4*>(dst) = v;
          }
        }
        __syncwarp();
      }
    }
    if (tri) item_stamp(debug, item_k, 4);
    ++item_k;
  }
}
