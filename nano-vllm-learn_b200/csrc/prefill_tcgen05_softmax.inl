// Body of the softmax warpgroups of prefill_tcgen05_kernel (textually included inside the kernel).
// Warpgroup t owns query tile t of every work item: thread `row` owns TMEM lane `row` of S_t / O_t.
{
  const int t = warp >> 2;  // 0 = tile A, 1 = tile B
  const int row = tid & 127;
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const uint32_t s_addr = tmem + lane_base + static_cast<uint32_t>(t * TN);
  const uint32_t o_addr = tmem + lane_base + 2u * TN + static_cast<uint32_t>(t * kHeadDim);
  const float c = p.scale_log2;
  uint32_t s_cnt = 0;  // S tiles consumed by this warpgroup (phase of s_full[t])
  uint32_t o_cnt = 0;  // O commits consumed (phase of o_full[t]); one per PV, waited exactly once each

  for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
    const Item it = decode_item(p, w, num_pairs, batch);
    const int nt = it.n_t[t];
    if (!it.live || nt == 0) continue;
    const int qrow = it.m0 + t * TM + row;              // row inside the sequence
    const int limit = min(it.shift + qrow, it.Lk - 1);  // last visible key of this row
    const bool dump = kDebug && debug != nullptr && it.m0 == 0 && it.h == 0 && it.seq == 0 && t == 0;
    float m_used = -INFINITY;  // max the stored P/O/l are relative to (may lag the true max by < 2^8)
    float l_run = 0.f;

    const bool tr = kDebug && debug != nullptr && blockIdx.x == 0 && w == static_cast<int>(blockIdx.x) && row == 0;
    for (int j = 0; j < nt; ++j) {
      const int t_base = j * TN;
      if (tr) trace_stamp(debug, 1 + t, j, 0);
      tc::mbar_wait(bar(kSFull + t), s_cnt & 1u);
      ++s_cnt;
      tc::fence_after_sync();
      if (tr) trace_stamp(debug, 1 + t, j, 1);
      uint32_t s[TN];
      if (dbg_mode & 2u) {
#pragma unroll
        for (int i = 0; i < TN; ++i) s[i] = 0u;
      } else {
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) tc::tmem_ld32p(s_addr + cc * 32, s + cc * 32);
        tc::tmem_ld_wait();
      }
      if (tr) trace_stamp(debug, 1 + t, j, 2);
      if (dump && j == 0) {
#pragma unroll
        for (int i = 0; i < TN; ++i) debug[row * TN + i] = __uint_as_float(s[i]);
      }
      const bool need_mask = (t_base + TN - 1 > it.shift + it.m0 + t * TM) || (t_base + TN > it.Lk);
      if (need_mask) {
#pragma unroll
        for (int i = 0; i < TN; ++i)
          if (t_base + i > limit) s[i] = 0xff800000u;  // -inf
      }
      // four independent max chains instead of one 128-long dependent chain
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int i = 0; i < TN; i += 4) {
        mx0 = fmaxf(mx0, __uint_as_float(s[i]));
        mx1 = fmaxf(mx1, __uint_as_float(s[i + 1]));
        mx2 = fmaxf(mx2, __uint_as_float(s[i + 2]));
        mx3 = fmaxf(mx3, __uint_as_float(s[i + 3]));
      }
      const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));

      if (j == 0) {
        m_used = mx;
      } else {
        // PV(j-1) finished before S(j) could be produced (in-order tensor pipe): this wait only keeps phases in step
        tc::mbar_wait(bar(kOFull + t), o_cnt & 1u);
        ++o_cnt;
        tc::fence_after_sync();
        const bool need = mx * c > m_used * c + kRescaleThreshold;
        if (__any_sync(0xffffffffu, need)) {
          const float alpha = need ? fast_exp2(m_used * c - mx * c) : 1.f;
          if (need) m_used = mx;
          l_run *= alpha;
#pragma unroll
          for (int cc = 0; cc < 4; ++cc) {
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
      const float m_sc = (m_used == -INFINITY) ? 0.f : m_used * c;
      float psum0 = 0.f, psum1 = 0.f;
      if (!(dbg_mode & 1u)) {
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float p0 = fast_exp2(fmaf(__uint_as_float(s[cc * 32 + 2 * i]), c, -m_sc));
            const float p1 = fast_exp2(fmaf(__uint_as_float(s[cc * 32 + 2 * i + 1]), c, -m_sc));
            psum0 += p0;
            psum1 += p1;
            pk[i] = pack_bf16x2(p0, p1);
          }
          tc::tmem_st16(s_addr + cc * 16, pk);
        }
      }
      if (tr) trace_stamp(debug, 1 + t, j, 4);
      tc::tmem_st_wait();
      tc::fence_before_sync();
      tc::mbar_arrive(bar(kPFull + t));
      if (tr) trace_stamp(debug, 1 + t, j, 5);
      l_run += psum0 + psum1;
    }

    // ---- epilogue: O_t (TMEM) / l -> bf16 -> global; rows beyond the sequence are not stored
    tc::mbar_wait(bar(kOFull + t), o_cnt & 1u);
    ++o_cnt;
    tc::fence_after_sync();
    const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
    uint4* dst =
        reinterpret_cast<uint4*>(p.out + (static_cast<int64_t>(it.q0 + qrow) * p.num_q_heads + it.h) * kHeadDim);
    uint32_t o[kHeadDim];
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) tc::tmem_ld32p(o_addr + cc * 32, o + cc * 32);
    tc::tmem_ld_wait();
    tc::fence_before_sync();
    tc::mbar_arrive(bar(kOEmpty + t));  // the next item's first PV may overwrite O_t now
    if (dump && nt == 1) {
#pragma unroll
      for (int i = 0; i < kHeadDim; ++i) debug[TM * TN + row * kHeadDim + i] = __uint_as_float(o[i]);
    }
    if (qrow < it.Lq) {
#pragma unroll
      for (int i = 0; i < kHeadDim / 8; ++i) {
        uint4 v;
        v.x = pack_bf16x2(__uint_as_float(o[8 * i + 0]) * inv, __uint_as_float(o[8 * i + 1]) * inv);
        v.y = pack_bf16x2(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv);
        v.z = pack_bf16x2(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv);
        v.w = pack_bf16x2(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv);
        dst[i] = v;
      }
    }
  }
}
