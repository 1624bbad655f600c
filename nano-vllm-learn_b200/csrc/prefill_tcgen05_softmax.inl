// Body of the softmax warpgroups of prefill_tcgen05_kernel (textually included inside the kernel).
// Warpgroup t owns query tile t: thread `row` owns TMEM lane `row` of S_t / O_t.
{
  const int t = warp >> 2;  // 0 = tile A, 1 = tile B
  const int nt = n_t[t];
  if (nt > 0) {
    const int row = tid & 127;
    const int qrow = m0 + t * TM + row;  // row inside the sequence
    const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
    const uint32_t s_addr = tmem + lane_base + static_cast<uint32_t>(t * TN);
    const uint32_t o_addr = tmem + lane_base + 2u * TN + static_cast<uint32_t>(t * kHeadDim);
    const int limit = min(shift + qrow, Lk - 1);  // last visible key of this row
    const float c = p.scale_log2;
    const bool dump = kDebug && debug != nullptr && mbp == 0 && h == 0 && seq == 0 && t == 0;
    float m_used = -INFINITY;  // max the stored P/O/l are relative to (may lag the true max by < 2^8)
    float l_run = 0.f;

    for (int j = 0; j < nt; ++j) {
      const int t_base = j * TN;
      tc::mbar_wait(bar(kSFull + t), static_cast<uint32_t>(j) & 1u);
      tc::fence_after_sync();
      uint32_t s[TN];
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) tc::tmem_ld32p(s_addr + cc * 32, s + cc * 32);
      tc::tmem_ld_wait();
      if (dump && j == 0) {
#pragma unroll
        for (int i = 0; i < TN; ++i) debug[row * TN + i] = __uint_as_float(s[i]);
      }
      const bool need_mask = (t_base + TN - 1 > shift + m0 + t * TM) || (t_base + TN > Lk);
      if (need_mask) {
#pragma unroll
        for (int i = 0; i < TN; ++i)
          if (t_base + i > limit) s[i] = 0xff800000u;  // -inf
      }
      float mx = -INFINITY;
#pragma unroll
      for (int i = 0; i < TN; ++i) mx = fmaxf(mx, __uint_as_float(s[i]));

      if (j == 0) {
        m_used = mx;
      } else {
        // PV(j-1) finished before S(j) could be produced (in-order tensor pipe): this wait only keeps phases in step
        tc::mbar_wait(bar(kOFull + t), static_cast<uint32_t>(j - 1) & 1u);
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
      const float m_sc = (m_used == -INFINITY) ? 0.f : m_used * c;
      float psum0 = 0.f, psum1 = 0.f;
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
      tc::tmem_st_wait();
      tc::fence_before_sync();
      tc::mbar_arrive(bar(kPFull + t));
      l_run += psum0 + psum1;
    }

    // ---- epilogue: O_t (TMEM) / l -> bf16 -> global; rows beyond the sequence are not stored
    tc::mbar_wait(bar(kOFull + t), static_cast<uint32_t>(nt - 1) & 1u);
    tc::fence_after_sync();
    const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
    uint4* dst = reinterpret_cast<uint4*>(p.out + (static_cast<int64_t>(q0 + qrow) * p.num_q_heads + h) * kHeadDim);
#pragma unroll
    for (int cc = 0; cc < 4; ++cc) {
      uint32_t r[32];
      tc::tmem_ld32(o_addr + cc * 32, r);
      tc::tmem_ld_wait();
      if (dump && nt == 1) {
#pragma unroll
        for (int i = 0; i < 32; ++i) debug[TM * TN + row * kHeadDim + cc * 32 + i] = __uint_as_float(r[i]);
      }
      if (qrow < Lq) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 v;
          v.x = pack_bf16x2(__uint_as_float(r[8 * i + 0]) * inv, __uint_as_float(r[8 * i + 1]) * inv);
          v.y = pack_bf16x2(__uint_as_float(r[8 * i + 2]) * inv, __uint_as_float(r[8 * i + 3]) * inv);
          v.z = pack_bf16x2(__uint_as_float(r[8 * i + 4]) * inv, __uint_as_float(r[8 * i + 5]) * inv);
          v.w = pack_bf16x2(__uint_as_float(r[8 * i + 6]) * inv, __uint_as_float(r[8 * i + 7]) * inv);
          dst[cc * 4 + i] = v;
        }
      }
    }
  }
}
