// EXPERIMENT (not part of the library; included by tools/rc_bench.cu only).  Range coder "version 2": exact division by a
// 65-bit magic and the renormalisation decision moved behind the multiply.  Byte-identical to version 1 and to the
// reference, but measured SLOWER on B200 (152 vs 137 cycles per symbol, profiles/r2a_rc_bench.txt): a lone warp pays
// about (1 + stall field) cycles per instruction (tools/sass_cost.py reproduces both measurements within 5 %), and this
// formulation needs 51 instructions per step against 41, which outweighs its shorter dependent chain.
#pragma once
#include "entropy.cuh"

namespace lbent {

// ---- range coder, version 2 (the per-block streams) -----------------------------------------------------------
// What bounds a lone warp is the DEPENDENT chain of a step, and in version 1 that chain ran through everything:
// multiply-high (34 cycles) -> shift -> three products -> xor -> mask -> select of (low, range) -> next multiply-high.
// Version 2 cuts it in two chains that run side by side:
//   * exact division by a 65-bit magic (Granlund-Montgomery): with ell = ceil(log2 d), mp = floor(2^64 (2^ell - d) / d) + 1,
//     floor(n / d) == (n * (2^64 + mp)) >> (64 + ell) for EVERY n < 2^64 -- no "increment" case, nothing to carry but n;
//   * the 0-or-1-byte renormalisation only multiplies the range by 2^8, and (rg << 8) * M == (rg * M) << 8: the 129-bit
//     product W = rg * (2^64 + mp) is formed from the UNSHIFTED range as soon as rg exists, and the renormalisation
//     decision (which needs low, the longer computation) only picks the bit window of W the quotient is read from:
//         q_next = W >> (56 + ell_next + 8 * zero)          zero = "no byte was emitted"
//     i.e. the decision enters one funnel shift before the next quotient instead of sitting in front of the multiply.
// Loop-carried state: (low, q) with q = range / total of the NEXT symbol.  Chain A: q -> rg = q * w -> W -> q'.  Chain B:
// q -> low + q * lo, low + q * (lo + w) -> xor -> zero -> shift amount.  Rare steps (two or more bytes, a range clamp)
// are detected exactly as in version 1 and the group is recoded with rc_step_gen.
LB_HD void div_magic65(uint32_t d, uint64_t &mp, uint32_t &ell)
{
#ifdef __CUDA_ARCH__
    ell = d > 1 ? 32u - (uint32_t)__clz(d - 1) : 0u;
#else
    ell = d > 1 ? 32u - (uint32_t)__builtin_clz(d - 1) : 0u;
#endif
    const uint64_t num = (ell >= 32 ? (1ull << 32) : (1ull << ell)) - d;      // < d
#ifdef __CUDA_ARCH__
    const double rcp = 1.0 / (double)d;
    uint64_t r1, r2;
    const uint64_t q1 = div_64_32(num << 32, d, rcp, r1);
    const uint64_t q0 = div_64_32(r1 << 32, d, rcp, r2);
    mp = ((q1 << 32) | q0) + 1;
#else
    mp = (uint64_t)((((unsigned __int128)num) << 64) / d) + 1;
#endif
}

// floor(n / d) through the 65-bit magic of d, any n
LB_HD uint64_t div65(uint64_t n, uint64_t mp, uint32_t ell)
{
#ifdef __CUDA_ARCH__
    const uint64_t t = __umul64hi(n, mp);
#else
    const uint64_t t = (uint64_t)(((unsigned __int128)n * mp) >> 64);
#endif
    const uint64_t v = t + n;
    const uint64_t c = v < t ? 1ull : 0ull;                  // bit 64 of the 65-bit sum
    if (ell == 0) return v;
    return (v >> ell) | (ell >= 64 ? 0ull : ((c << 1) << (63 - ell)));
}

// per-symbol record of version 2: (lo, w, mp lo, mp hi) + byte ell
LB_D void store_symrec65(uint4 *__restrict__ rec, uint8_t *__restrict__ sh, uint64_t pos, uint32_t lo, uint32_t w, uint32_t tot)
{
    uint64_t mp; uint32_t ell;
    div_magic65(tot, mp, ell);
    rec[pos] = make_uint4(lo, w, (uint32_t)mp, (uint32_t)(mp >> 32));
    sh[pos] = (uint8_t)ell;
}

// Generic step: any symbol, any state.  (low, q) -> (low', q') where q' = range' / total of the next symbol, whose magic
// is (mp_n, ell_n).  Bytes are appended through `put`.
template <typename PUT>
LB_HD void rc_step_gen(uint64_t &low, uint64_t &q, uint32_t lo, uint32_t w, uint64_t mp_n, uint32_t ell_n, PUT &put)
{
    low += (uint64_t)lo * q;
    uint64_t range = q * (uint64_t)w;
    for (int guard = 0; guard < (int)RC_MAX_BYTES_PER_SYMBOL; guard++) {
        const uint32_t lh = (uint32_t)(low >> 32);
        if ((lh ^ (uint32_t)((low + range) >> 32)) >= (1u << 24)) {
            if ((uint32_t)(range >> 32) >= (1u << 16)) break;
            range = (0 - low) & (RC_BOTTOM - 1);
        }
        put((uint8_t)(lh >> 24));
        range <<= 8;
        low <<= 8;
    }
    q = div65(range, mp_n, ell_n);
}

LB_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t amt)       // bits [amt, amt + 32) of hi:lo, amt <= 31
{
#ifdef __CUDA_ARCH__
    return __funnelshift_r(lo, hi, amt);
#else
    return (uint32_t)((((uint64_t)hi << 32) | lo) >> amt);
#endif
}

// Common-case step (0 or 1 byte, no clamp), straight-line.  rec = (lo, w, lo + w, -), next magic (mp_n, ell_n <= 23).
// The byte is stored unconditionally at out[on]; on advances by 0 or 1.  rmin accumulates the rarity test.
template <typename STB>
LB_HD void rc_step_fast(uint64_t &low, uint64_t &q, uint32_t &on, uint32_t &rmin, uint32_t lo, uint32_t w, uint32_t low_w,
                        uint32_t mpl, uint32_t mph, uint32_t ell_n, STB &stb)
{
    const uint32_t ql = (uint32_t)q, qh = (uint32_t)(q >> 32);
    // chain A: rg = q * w (fits: q * total <= range)
    const uint64_t rg0 = (uint64_t)ql * w;
    const uint32_t rl = (uint32_t)rg0, rh = (uint32_t)(rg0 >> 32) + qh * w;
    // chain B: lw = low + q * lo, top = low + q * (lo + w)
    const uint64_t lw = low + (uint64_t)ql * lo + ((uint64_t)(qh * lo) << 32);
    const uint64_t tp = low + (uint64_t)ql * low_w + ((uint64_t)(qh * low_w) << 32);
    const uint32_t lh = (uint32_t)(lw >> 32), ll = (uint32_t)lw;
    const uint32_t x = lh ^ (uint32_t)(tp >> 32);
    const bool zero = x >= (1u << 24);
    // W = rg * (2^64 + mp): words y1 (bits 32..63) .. y4 (bit 128)
    const uint64_t t0 = (uint64_t)rl * mpl, t1 = (uint64_t)rh * mpl, t2 = (uint64_t)rl * mph;
    const uint64_t t3 = (uint64_t)rh * mph + rl;                                           // + rg << 64, low word
    const uint64_t s1 = (t0 >> 32) + (uint32_t)t1 + (uint32_t)t2;                          // word 1 + carries (< 3 * 2^32)
    const uint64_t s2 = (s1 >> 32) + (t1 >> 32) + (t2 >> 32) + (uint32_t)t3;               // word 2 + carries
    const uint64_t s3 = (s2 >> 32) + (t3 >> 32) + rh;                                      // word 3 + carry = bit 128
    const uint32_t y1 = (uint32_t)s1, y2 = (uint32_t)s2, y3 = (uint32_t)s3, y4 = (uint32_t)(s3 >> 32);
    const uint32_t v0 = funnel_r(y1, y2, 24), v1 = funnel_r(y2, y3, 24), v2 = funnel_r(y3, y4, 24);
    const uint32_t amt = zero ? ell_n + 8u : ell_n;
    const uint32_t nql = funnel_r(v0, v1, amt), nqh = funnel_r(v1, v2, amt);
    // byte out, state
    stb(on, lh >> 24);
    on += zero ? 0u : 1u;
    const uint32_t r1h = (rh << 8) | (rl >> 24);
    rmin = min(rmin, min(x, zero ? rh : r1h));      // rare: second byte settled too (x < 2^16) or range' < 2^48
    low = zero ? lw : (lw << 8);
    (void)ll;
    q = ((uint64_t)nqh << 32) | nql;
}

constexpr int RC2_BATCH = 64;
constexpr int RC2_OBUF = RC2_BATCH * RC_MAX_BYTES_PER_SYMBOL + 32;
constexpr uint32_t RC2_MAX_ELL = 23;     // fast step: ell + 8 must stay a 5-bit shift amount

struct Sts8Shared {
    LB_D void operator()(uint32_t addr, uint32_t v) const { sts8(addr, v); }
};
struct PutShared {
    uint32_t &on;
    LB_D void operator()(uint8_t b) { sts8(on++, b); }
};

// RangeEncoder::encode / flush, C/RangeCoder.cpp:106-145.  One warp per stream; lane 0 runs the recurrence, the other
// lanes stage the records (unpacked into the form lane 0 consumes) and copy the output bytes.
__global__ void __launch_bounds__(32) k_range_encode2(const uint4 *__restrict__ rec_all, const uint8_t *__restrict__ sh_all,
                                                      const StreamDesc *__restrict__ descs, uint8_t *__restrict__ out_all,
                                                      uint64_t *__restrict__ out_size, int always_flush, uint32_t *overflow)
{
    __shared__ uint4 sa[2][RC2_BATCH + 4];        // (lo, w, lo + w, ell)
    __shared__ uint2 sm[2][RC2_BATCH + 4];        // mp
    __shared__ uint8_t sout[RC2_OBUF];
    const int lane = threadIdx.x;
    const StreamDesc d = descs[blockIdx.x];
    const uint4 *rec = rec_all + d.rec_base;
    const uint8_t *sh = sh_all + d.rec_base;
    uint8_t *const out0 = out_all + d.out_base;
    uint64_t opos = 0;
    uint64_t low = 0, q = 0;                      // lane 0
    // identity symbol (total 1: mp = 1, ell = 0 gives q = range; lo = 0, w = 1): pads the last batch
    const uint4 idle = make_uint4(0u, 1u, 1u, 0u);
    uint4 na = idle, nb = idle;
    uint32_t nsa = 0u, nsb = 0u;
    if ((uint64_t)lane < d.nsym) { na = rec[lane]; nsa = sh[lane]; }
    if ((uint64_t)lane + 32 < d.nsym) { nb = rec[lane + 32]; nsb = sh[lane + 32]; }
    if (lane == 0) q = div65(~0ull, ((uint64_t)na.w << 32) | na.z, nsa);      // range = 2^64 - 1 over the first total
    for (uint64_t base = 0; base < d.nsym; base += RC2_BATCH) {
        const int buf = (int)((base / RC2_BATCH) & 1);
        sa[buf][lane] = make_uint4(na.x, na.y, na.x + na.y, nsa);
        sa[buf][lane + 32] = make_uint4(nb.x, nb.y, nb.x + nb.y, nsb);
        sm[buf][lane] = make_uint2(na.z, na.w);
        sm[buf][lane + 32] = make_uint2(nb.z, nb.w);
        uint32_t wide = (nsa > RC2_MAX_ELL) || (nsb > RC2_MAX_ELL);
        const uint64_t in = base + RC2_BATCH + lane;
        na = idle; nb = idle; nsa = 0u; nsb = 0u;
        if (in < d.nsym) { na = rec[in]; nsa = sh[in]; }
        if (in + 32 < d.nsym) { nb = rec[in + 32]; nsb = sh[in + 32]; }
        if (lane == 0) {      // the slot past the batch = first record of the next batch (its magic ends this batch's last step)
            sa[buf][RC2_BATCH] = make_uint4(na.x, na.y, na.x + na.y, nsa);
            sm[buf][RC2_BATCH] = make_uint2(na.z, na.w);
            wide |= nsa > RC2_MAX_ELL;
        }
        wide = __any_sync(0xffffffffu, wide);
        __syncwarp();
        const uint32_t sout_a = (uint32_t)__cvta_generic_to_shared(sout);
        uint32_t on = sout_a;
        if (lane == 0) {
            const uint32_t pa = (uint32_t)__cvta_generic_to_shared(sa[buf]);      // record j at pa + 16 j
            const uint32_t pm = (uint32_t)__cvta_generic_to_shared(sm[buf]);      // magic of j at pm + 8 j
            Sts8Shared stb;
            if (wide) {
#pragma unroll 1
                for (int j = 0; j < RC2_BATCH; j++) {
                    const uint4 a = lds128(pa + 16 * j), an = lds128(pa + 16 * j + 16);
                    const uint2 mn = lds64(pm + 8 * j + 8);
                    PutShared put{on};
                    rc_step_gen(low, q, a.x, a.y, ((uint64_t)mn.y << 32) | mn.x, an.w, put);
                }
            } else {
                uint4 a0 = lds128(pa);
#pragma unroll 1
                for (int j = 0; j < RC2_BATCH; j += 4) {
                    const uint4 a1 = lds128(pa + 16 * j + 16), a2 = lds128(pa + 16 * j + 32), a3 = lds128(pa + 16 * j + 48);
                    const uint4 a4 = lds128(pa + 16 * j + 64);
                    const uint2 m1 = lds64(pm + 8 * j + 8), m2 = lds64(pm + 8 * j + 16), m3 = lds64(pm + 8 * j + 24);
                    const uint2 m4 = lds64(pm + 8 * j + 32);
                    const uint64_t low_g = low, q_g = q;
                    const uint32_t on_g = on;
                    uint32_t rmin = 0xffffffffu;
                    rc_step_fast(low, q, on, rmin, a0.x, a0.y, a0.z, m1.x, m1.y, a1.w, stb);
                    rc_step_fast(low, q, on, rmin, a1.x, a1.y, a1.z, m2.x, m2.y, a2.w, stb);
                    rc_step_fast(low, q, on, rmin, a2.x, a2.y, a2.z, m3.x, m3.y, a3.w, stb);
                    rc_step_fast(low, q, on, rmin, a3.x, a3.y, a3.z, m4.x, m4.y, a4.w, stb);
                    if (__builtin_expect(rmin < (1u << 16), 0)) {
                        low = low_g; q = q_g; on = on_g;
                        PutShared put{on};
                        rc_step_gen(low, q, a0.x, a0.y, ((uint64_t)m1.y << 32) | m1.x, a1.w, put);
                        rc_step_gen(low, q, a1.x, a1.y, ((uint64_t)m2.y << 32) | m2.x, a2.w, put);
                        rc_step_gen(low, q, a2.x, a2.y, ((uint64_t)m3.y << 32) | m3.x, a3.w, put);
                        rc_step_gen(low, q, a3.x, a3.y, ((uint64_t)m4.y << 32) | m4.x, a4.w, put);
                    }
                    a0 = a4;
                }
            }
        }
        on -= sout_a;
        on = __shfl_sync(0xffffffffu, on, 0);
        for (uint32_t i = lane; i < on; i += 32)
            if (opos + i < d.out_cap) out0[opos + i] = sout[i];
        opos += on;
        __syncwarp();
    }
    if (lane == 0) {
        uint64_t nout = opos;
        if (always_flush || nout > 0) {                       // DnaEncoder::writeBlock flushes only a non-empty buffer
            for (int x = 0; x < 8; x++) {
                if (nout < d.out_cap) out0[nout] = (uint8_t)(low >> 56);
                nout++;
                low <<= 8;
            }
        }
        out_size[blockIdx.x] = nout;
        if (nout > d.out_cap) *overflow = 1;
    }
}


}  // namespace lbent
